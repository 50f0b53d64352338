// fwd kernels of the architectures in FB_SPEC_LIST_d (fb_spec_lists.h)
#include "fb_spec_ops_fwd.cuh"
#include "fb_spec_lists.h"
static const FbSpecFwdEntry entries[] = {FB_SPEC_LIST_d(FB_SPEC_FWD_ENTRY)};
const FbSpecFwdTable fb_spec_fwd_table_d = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
