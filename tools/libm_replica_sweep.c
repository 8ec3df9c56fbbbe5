// libm_replica_sweep.c — exhaustive check of the library's host builds of glibc's expf / log1pf / logf / log10f
// (fast-ctc-decode_b200/csrc/phred.cuh, exported as fctc_host_*) against the system libm, over ALL float bit
// patterns of each function's domain (NaN results count as equal whatever their payload).  tests/test_cabi_cpu.py
// holds strided versions of the same sweeps; this is the full one (about a minute on 8 cores).
//
//   gcc -O2 -o /tmp/sweep tools/libm_replica_sweep.c -lm -lpthread -ldl
//   /tmp/sweep fast-ctc-decode_b200/libfctc_b200.so
#define _GNU_SOURCE
#include <dlfcn.h>
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef void (*batch_fn)(const float*, float*, int64_t);
typedef struct { uint64_t lo, hi; batch_fn mine; float (*ref)(float); uint64_t bad; uint32_t first; } job_t;

static void* run(void* a) {
    job_t* j = a;
    enum { B = 1 << 16 };
    float* x = malloc(B * 4), *y = malloc(B * 4);
    j->bad = 0;
    for (uint64_t u0 = j->lo; u0 < j->hi; u0 += B) {
        const uint64_t n = j->hi - u0 < B ? j->hi - u0 : B;
        for (uint64_t i = 0; i < n; ++i) { uint32_t u = (uint32_t)(u0 + i); memcpy(&x[i], &u, 4); }
        j->mine(x, y, (int64_t)n);
        for (uint64_t i = 0; i < n; ++i) {
            const float g = j->ref(x[i]);
            uint32_t a_, b_;
            memcpy(&a_, &y[i], 4); memcpy(&b_, &g, 4);
            if (a_ != b_ && !(y[i] != y[i] && g != g)) { if (!j->bad) j->first = (uint32_t)(u0 + i); j->bad++; }
        }
    }
    free(x); free(y);
    return 0;
}

int main(int argc, char** argv) {
    void* lib = dlopen(argc > 1 ? argv[1] : "fast-ctc-decode_b200/libfctc_b200.so", RTLD_NOW);
    if (!lib) { fprintf(stderr, "%s\n", dlerror()); return 2; }
    const char* names[6] = {"fctc_host_expf", "fctc_host_log1pf", "fctc_host_logf", "fctc_host_log10f",
                            "fctc_host_expf_nonpos", "fctc_host_log1pf_unit"};
    float (*refs[6])(float) = {expf, log1pf, logf, log10f, expf, log1pf};
    const int nt = argc > 2 ? atoi(argv[2]) : 8;
    int rc = 0;
    for (int f = 0; f < 6; ++f) {
        batch_fn fn = (batch_fn)dlsym(lib, names[f]);
        if (!fn) { fprintf(stderr, "missing %s\n", names[f]); return 2; }
        uint64_t lo = 0, hi = 1ull << 32;
        if (f == 3) { lo = 0x00800000ull; hi = 0x7fc00001ull; }  // log10f: positive normal numbers, +inf, NaN (phred's domain)
        if (f == 4) { lo = 0x80000000ull; hi = 1ull << 32; }      // -0, every negative number, -inf, the negative NaNs
        if (f == 5) { lo = 0; hi = 0x3f800001ull; }               // [0, 1]
        pthread_t th[64]; job_t jobs[64];
        for (int i = 0; i < nt; ++i) {
            jobs[i].lo = lo + (hi - lo) * i / nt; jobs[i].hi = lo + (hi - lo) * (i + 1) / nt;
            jobs[i].mine = fn; jobs[i].ref = refs[f];
            pthread_create(&th[i], 0, run, &jobs[i]);
        }
        uint64_t bad = 0;
        for (int i = 0; i < nt; ++i) {
            pthread_join(th[i], 0);
            bad += jobs[i].bad;
            if (jobs[i].bad) printf("  first difference in part %d: bit pattern 0x%08x\n", i, jobs[i].first);
        }
        printf("%s vs libm: %llu differences over %llu bit patterns\n", names[f], (unsigned long long)bad,
               (unsigned long long)(hi - lo));
        if (bad) rc = 1;
        if (f >= 4) {  // the remaining members of the restricted domains: +0 and the positive NaNs
            float xs[4], ys[4];
            const uint32_t pat[4] = {0x00000000u, 0x7fc00000u, 0x7f800001u, 0x7fffffffu};
            for (int i = 0; i < 4; ++i) memcpy(&xs[i], &pat[i], 4);
            fn(xs, ys, 4);
            for (int i = 0; i < 4; ++i) {
                const float g = refs[f](xs[i]);
                if (memcmp(&ys[i], &g, 4) && !(ys[i] != ys[i] && g != g)) { printf("  %s differs at 0x%08x\n", names[f], pat[i]); rc = 1; }
            }
        }
    }
    return rc;
}
