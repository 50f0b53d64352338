// Two/three-hidden-layer mid-size architectures (BM-2, BM-9..13, CP-2..11, DF-*).
#include "fb_spec_ops.cuh"
static const FbSpecEntry entries[] = {
    FB_SPEC_ENTRY(16, 32, 16, 1), FB_SPEC_ENTRY(16, 30, 20, 1), FB_SPEC_ENTRY(12, 32, 32, 1),
    FB_SPEC_ENTRY(6, 32, 32, 1),  FB_SPEC_ENTRY(30, 16, 16, 16, 1),
};
const FbSpecTable fb_spec_table_d = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
