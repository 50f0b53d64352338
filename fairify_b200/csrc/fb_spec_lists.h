// fb_spec_lists.h -- the architectures that get compile-time-specialised kernels, grouped into build units.
// Every unit is compiled three times (forward / profile / bounds kernels: fb_specF_x.cu, fb_specP_x.cu, fb_specI_x.cu).
#pragma once
// Adult-census architectures (AC-1, AC-3, AC-6, AC-8..AC-12)
#define FB_SPEC_LIST_a(X) X(13, 16, 8, 1) X(13, 50, 1) X(13, 12, 12, 1) X(13, 5, 5, 1) X(13, 3, 3, 3, 3, 1) \
    X(13, 5, 5, 5, 5, 1) X(13, 10, 10, 10, 10, 1) X(13, 5, 5, 5, 5, 5, 5, 5, 5, 5, 1)
// one-hidden-layer, 100-wide architectures (AC-2, GC-2, BM-3)
#define FB_SPEC_LIST_b(X) X(13, 100, 1) X(20, 100, 1) X(16, 100, 1)
// small German-credit / bank / compas architectures (GC-1, GC-3, GC-4, BM-5, BM-6, CP-1)
#define FB_SPEC_LIST_c(X) X(20, 50, 1) X(20, 9, 1) X(20, 6, 4, 1) X(16, 22, 10, 1) X(16, 9, 9, 1) X(6, 16, 8, 1)
// two/three-hidden-layer mid-size architectures (BM-2, BM-9..13, CP-2..11, DF-*)
#define FB_SPEC_LIST_d(X) X(16, 32, 16, 1) X(16, 30, 20, 1) X(12, 32, 32, 1) X(6, 32, 32, 1) X(30, 16, 16, 16, 1)
// BM-1 (16-64-16-1, 2064 weights)
#define FB_SPEC_LIST_e(X) X(16, 64, 16, 1)
// deep funnel architectures (AC-7 13-64-32-16-8-4-1, BM-8 16-64-32-16-8-4-1): ~3.6k weights each
#define FB_SPEC_LIST_f(X) X(13, 64, 32, 16, 8, 4, 1) X(16, 64, 32, 16, 8, 4, 1)
// GC-5 (20-64-32-16-8-4-1, 4004 weights): fits the kernel-parameter space with float32-stored biases
#define FB_SPEC_LIST_g(X) X(20, 64, 32, 16, 8, 4, 1)

// Architectures whose activation profile runs on the FP64 matrix instruction with compile-time widths (fb_specM_x.cu):
// the four wide reference networks, and the mid-size ones whose hidden-to-hidden layers hold most of the weights (for the
// rest -- one hidden layer, or layers far from a multiple of 8 neurons -- the padding of the 8-column tiles and the
// one-column logit tile cost more than the matrix instruction saves; they stay on fb_profile_spec_kernel)
#define FB_SPEC_MMA_LIST_a(X) X(13, 100, 100, 1) X(13, 64, 64, 1)       /* AC-4, AC-5 */
#define FB_SPEC_MMA_LIST_b(X) X(16, 150, 100, 50, 1)                    /* BM-4 */
#define FB_SPEC_MMA_LIST_c(X) X(16, 64, 64, 1)                          /* BM-7 */
#define FB_SPEC_MMA_LIST_d(X) X(16, 64, 16, 1) X(13, 64, 32, 16, 8, 4, 1) X(16, 64, 32, 16, 8, 4, 1) X(20, 64, 32, 16, 8, 4, 1) /* BM-1, AC-7, BM-8, GC-5 */
// (measured and left on the FMA kernel: BM-2 16-32-16-1 70 % vs 53 % of FP64 peak with DMMA, BM-9 16-30-20-1 69 vs 49)
