/* orbx test/measurement probes (exported by liborbx.so; not part of the reference surface). */
#ifndef ORBX_TEST_H
#define ORBX_TEST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
/* device port of glibc-2.39 sinf/cosf on `count` consecutive float bit patterns starting at first_bits */
int orbx_test_sincosf(uint32_t first_bits, uint32_t count, float* h_sin, float* h_cos);
/* device cv::fastAtan2 on n (y,x) pairs */
int orbx_test_fast_atan2(const float* h_yx, int n, float* h_out);
/* POPC32 throughput of the whole GPU (8 independent chains per thread), POPC/s */
int orbx_test_popc_peak(int device, int iters, double* popc_per_s, double* ms);
#ifdef __cplusplus
}
#endif
#endif
