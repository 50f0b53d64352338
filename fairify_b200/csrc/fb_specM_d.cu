// DMMA profile kernels of the architectures in FB_SPEC_MMA_LIST_d (fb_spec_lists.h)
#define FB_MMA_SPEC_MS_MASK 7
#define FB_MMA_SPEC_DEFAULT_MS 4
#include "fb_spec_ops_mma.cuh"
#include "fb_spec_lists.h"
static const FbSpecMmaEntry entries[] = {FB_SPEC_MMA_LIST_d(FB_SPEC_MMA_ENTRY)};
const FbSpecMmaTable fb_spec_mma_table_d = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
