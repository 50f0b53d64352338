// ibp kernels of the architectures in FB_SPEC_LIST_e (fb_spec_lists.h)
#include "fb_spec_ops_ibp.cuh"
#include "fb_spec_lists.h"
static const FbSpecIbpEntry entries[] = {FB_SPEC_LIST_e(FB_SPEC_IBP_ENTRY)};
const FbSpecIbpTable fb_spec_ibp_table_e = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
