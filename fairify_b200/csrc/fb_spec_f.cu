// Deep funnel architectures (AC-7 13-64-32-16-8-4-1, BM-8 16-64-32-16-8-4-1): ~3.6k weights each, the largest
// parameter structs that still fit the 32 KB kernel-parameter space.
#include "fb_spec_ops.cuh"
static const FbSpecEntry entries[] = {
    FB_SPEC_ENTRY(13, 64, 32, 16, 8, 4, 1), FB_SPEC_ENTRY(16, 64, 32, 16, 8, 4, 1),
};
const FbSpecTable fb_spec_table_f = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
