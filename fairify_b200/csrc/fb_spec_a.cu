// Adult-census architectures (AC-1, AC-3, AC-6, AC-8..AC-12).
#include "fb_spec_ops.cuh"
static const FbSpecEntry entries[] = {
    FB_SPEC_ENTRY(13, 16, 8, 1),  FB_SPEC_ENTRY(13, 50, 1),          FB_SPEC_ENTRY(13, 12, 12, 1),
    FB_SPEC_ENTRY(13, 5, 5, 1),   FB_SPEC_ENTRY(13, 3, 3, 3, 3, 1),  FB_SPEC_ENTRY(13, 5, 5, 5, 5, 1),
    FB_SPEC_ENTRY(13, 10, 10, 10, 10, 1), FB_SPEC_ENTRY(13, 5, 5, 5, 5, 5, 5, 5, 5, 5, 1),
};
const FbSpecTable fb_spec_table_a = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
