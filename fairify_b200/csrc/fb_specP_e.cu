// profile kernels of the architectures in FB_SPEC_LIST_e (fb_spec_lists.h)
#include "fb_spec_ops_profile.cuh"
#include "fb_spec_lists.h"
static const FbSpecProfileEntry entries[] = {FB_SPEC_LIST_e(FB_SPEC_PROFILE_ENTRY)};
const FbSpecProfileTable fb_spec_profile_table_e = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
