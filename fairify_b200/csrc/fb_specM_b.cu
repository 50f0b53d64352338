// DMMA profile kernels of the architectures in FB_SPEC_MMA_LIST_b (fb_spec_lists.h)
#define FB_MMA_SPEC_MS_MASK 1
#define FB_MMA_SPEC_DEFAULT_MS 1
#include "fb_spec_ops_mma.cuh"
#include "fb_spec_lists.h"
static const FbSpecMmaEntry entries[] = {FB_SPEC_MMA_LIST_b(FB_SPEC_MMA_ENTRY)};
const FbSpecMmaTable fb_spec_mma_table_b = {entries, (int)(sizeof(entries) / sizeof(entries[0]))};
