// Small German-credit / bank / compas architectures (GC-1, GC-3, GC-4, BM-5, BM-6, CP-1).
#include "fb_spec_ops.cuh"
static const FbSpecEntry entries[] = {
    FB_SPEC_ENTRY(20, 50, 1),    FB_SPEC_ENTRY(20, 9, 1),    FB_SPEC_ENTRY(20, 6, 4, 1),
    FB_SPEC_ENTRY(16, 22, 10, 1), FB_SPEC_ENTRY(16, 9, 9, 1), FB_SPEC_ENTRY(6, 16, 8, 1),
};
const FbSpecTable fb_spec_table_c = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
