// One-hidden-layer, 100-wide architectures (AC-2, GC-2, BM-3).
#include "fb_spec_ops.cuh"
static const FbSpecEntry entries[] = {
    FB_SPEC_ENTRY(13, 100, 1), FB_SPEC_ENTRY(20, 100, 1), FB_SPEC_ENTRY(16, 100, 1),
};
const FbSpecTable fb_spec_table_b = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
