/* TEST INFRASTRUCTURE ONLY. Exhaustive pin of ref_sincosf.h against the host libm:
 * every float in [lo, hi] (default 0 .. 6.5), threads = argv[3]. Prints mismatch counts. */
#include <stdio.h>
#include <stdlib.h>
#include <pthread.h>
#include <gnu/libc-version.h>
#include "ref_sincosf.h"
typedef struct { uint32_t a, b; unsigned long bad_s, bad_c; } job_t;
static void* run(void* p)
{
    job_t* j = (job_t*)p;
    for (uint32_t u = j->a; u < j->b; ++u) {
        float f; memcpy(&f, &u, 4);
        float s = sinf(f), c = cosf(f), rs = ref_sinf(f), rc = ref_cosf(f);
        if (memcmp(&s, &rs, 4)) j->bad_s++;
        if (memcmp(&c, &rc, 4)) j->bad_c++;
    }
    return 0;
}
int main(int argc, char** argv)
{
    float lo = argc > 1 ? (float)atof(argv[1]) : 0.f, hi = argc > 2 ? (float)atof(argv[2]) : 6.5f;
    int nt = argc > 3 ? atoi(argv[3]) : 8;
    uint32_t ua, ub; memcpy(&ua, &lo, 4); memcpy(&ub, &hi, 4); ub += 1;
    pthread_t th[64]; job_t jb[64];
    uint64_t span = (uint64_t)ub - ua;
    for (int t = 0; t < nt; ++t) {
        jb[t].a = ua + (uint32_t)(span * t / nt); jb[t].b = ua + (uint32_t)(span * (t + 1) / nt);
        jb[t].bad_s = jb[t].bad_c = 0;
        pthread_create(&th[t], 0, run, &jb[t]);
    }
    unsigned long bs = 0, bc = 0;
    for (int t = 0; t < nt; ++t) { pthread_join(th[t], 0); bs += jb[t].bad_s; bc += jb[t].bad_c; }
    printf("{\"glibc\":\"%s\",\"n\":%lu,\"bad_sin\":%lu,\"bad_cos\":%lu}\n", gnu_get_libc_version(), (unsigned long)span, bs, bc);
    return (bs || bc) ? 1 : 0;
}
