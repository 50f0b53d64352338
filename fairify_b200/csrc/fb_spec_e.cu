// BM-1 (16-64-16-1, 2064 weights: the largest fully unrolled architecture).
#include "fb_spec_ops.cuh"
static const FbSpecEntry entries[] = {
    FB_SPEC_ENTRY(16, 64, 16, 1),
};
const FbSpecTable fb_spec_table_e = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
