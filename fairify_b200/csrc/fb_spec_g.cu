// GC-5 (20-64-32-16-8-4-1, 4004 weights): fits the kernel-parameter space with float32-stored biases.
#include "fb_spec_ops.cuh"
static const FbSpecEntry entries[] = {
    FB_SPEC_ENTRY(20, 64, 32, 16, 8, 4, 1),
};
const FbSpecTable fb_spec_table_g = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
