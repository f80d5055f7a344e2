/* oracle/gmo_pattern.c — TEST INFRASTRUCTURE (see gmo_pattern.h). Plain C, scalar loops. */
#define _GNU_SOURCE
#include "gmo_pattern.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define GMO_GAMMA 0x9E3779B97F4A7C15ull

/* SplitMix64 output function (Vigna, splitmix64.c, public domain). */
static uint64_t gmo_mix(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

uint64_t gmo_splitmix64_next(uint64_t* state) {
    *state += GMO_GAMMA;
    return gmo_mix(*state);
}

int gmo_lane_words(int pattern, uint64_t seed, uint64_t lane, uint64_t* w0, uint64_t* w1) {
    switch (pattern) {
        case GMO_PAT_SPLITMIX: {
            /* the lane-th output of the stream seeded with `seed`: state after lane+1 steps */
            uint64_t a = gmo_mix(seed + (lane + 1) * GMO_GAMMA);
            uint64_t swapped = (a << 32) | (a >> 32);
            *w0 = a;
            *w1 = ~swapped ^ lane;
            return 0;
        }
        case GMO_PAT_ZERO:
            *w0 = *w1 = 0;
            return 0;
        case GMO_PAT_ONES:
            *w0 = *w1 = ~(uint64_t)0;
            return 0;
        case GMO_PAT_CHECKER: {
            uint64_t v = (lane & 1) ? 0x5555555555555555ull : 0xAAAAAAAAAAAAAAAAull;
            if (seed & 1) v = ~v;
            *w0 = *w1 = v;
            return 0;
        }
        case GMO_PAT_ADDRESS:
            *w0 = seed + 16 * lane;
            *w1 = seed + 16 * lane + 8;
            return 0;
        default:
            return -1;
    }
}

static void put_le64(unsigned char* p, uint64_t v, size_t n) {
    for (size_t i = 0; i < n; ++i) p[i] = (unsigned char)(v >> (8 * i));
}

static uint64_t get_le64(const unsigned char* p, size_t n) {
    uint64_t v = 0;
    for (size_t i = 0; i < n; ++i) v |= (uint64_t)p[i] << (8 * i);
    return v;
}

static int fill_range(unsigned char* base, size_t bytes, uint64_t lane_begin, uint64_t lane_end, int pattern, uint64_t seed) {
    for (uint64_t L = lane_begin; L < lane_end; ++L) {
        uint64_t w0, w1;
        if (gmo_lane_words(pattern, seed, L, &w0, &w1) != 0) return -1;
        size_t off = (size_t)L * 16;
        size_t left = bytes - off;
        put_le64(base + off, w0, left >= 8 ? 8 : left);
        if (left > 8) put_le64(base + off + 8, w1, left >= 16 ? 8 : left - 8);
    }
    return 0;
}

int gmo_pattern_fill(void* buf, size_t bytes, int pattern, uint64_t seed) {
    if (pattern < 0 || pattern >= GMO_PAT_COUNT) return -1;
    uint64_t lanes = (bytes + 15) / 16;
    return fill_range((unsigned char*)buf, bytes, 0, lanes, pattern, seed);
}

static int verify_range(const unsigned char* base, size_t bytes, uint64_t lane_begin, uint64_t lane_end, int pattern,
                        uint64_t seed, gmo_verify_result* r) {
    r->mismatch_lanes = 0;
    r->first_bad_lane = UINT64_MAX;
    r->xor64 = 0;
    r->sum64 = 0;
    for (uint64_t L = lane_begin; L < lane_end; ++L) {
        uint64_t e0, e1;
        if (gmo_lane_words(pattern, seed, L, &e0, &e1) != 0) return -1;
        size_t off = (size_t)L * 16;
        size_t left = bytes - off;
        size_t n0 = left >= 8 ? 8 : left;
        size_t n1 = left > 8 ? (left >= 16 ? 8 : left - 8) : 0;
        uint64_t g0 = get_le64(base + off, n0);
        uint64_t g1 = n1 ? get_le64(base + off + 8, n1) : 0;
        uint64_t m0 = n0 == 8 ? ~(uint64_t)0 : (((uint64_t)1 << (8 * n0)) - 1);
        uint64_t m1 = n1 == 8 ? ~(uint64_t)0 : (n1 ? (((uint64_t)1 << (8 * n1)) - 1) : 0);
        r->xor64 ^= g0 ^ g1;
        r->sum64 += g0 + g1;
        if (((g0 ^ e0) & m0) || ((g1 ^ e1) & m1)) {
            if (r->mismatch_lanes == 0) r->first_bad_lane = L;
            r->mismatch_lanes++;
        }
    }
    return 0;
}

int gmo_pattern_verify(const void* buf, size_t bytes, int pattern, uint64_t seed, gmo_verify_result* out) {
    if (pattern < 0 || pattern >= GMO_PAT_COUNT || out == NULL) return -1;
    uint64_t lanes = (bytes + 15) / 16;
    return verify_range((const unsigned char*)buf, bytes, 0, lanes, pattern, seed, out);
}

int gmo_pattern_checksum(size_t bytes, int pattern, uint64_t seed, uint64_t* xor64, uint64_t* sum64) {
    if (pattern < 0 || pattern >= GMO_PAT_COUNT) return -1;
    uint64_t x = 0, s = 0;
    uint64_t full = bytes / 16;
    for (uint64_t L = 0; L < full; ++L) {
        uint64_t w0, w1;
        gmo_lane_words(pattern, seed, L, &w0, &w1);
        x ^= w0 ^ w1;
        s += w0 + w1;
    }
    size_t tail = bytes % 16;
    if (tail) {
        uint64_t w0, w1;
        unsigned char tmp[16];
        gmo_lane_words(pattern, seed, full, &w0, &w1);
        put_le64(tmp, w0, 8);
        put_le64(tmp + 8, w1, 8);
        uint64_t g0 = get_le64(tmp, tail >= 8 ? 8 : tail);
        uint64_t g1 = tail > 8 ? get_le64(tmp + 8, tail - 8) : 0;
        x ^= g0 ^ g1;
        s += g0 + g1;
    }
    if (xor64) *xor64 = x;
    if (sum64) *sum64 = s;
    return 0;
}

/* ---- threaded CPU baseline ---------------------------------------------------- */
typedef struct {
    unsigned char* a;
    unsigned char* b;
    size_t bytes;
    uint64_t lane_begin, lane_end;
    uint64_t seed;
    int phase;
    gmo_verify_result res;
} worker_arg;

static void* worker(void* p) {
    worker_arg* w = (worker_arg*)p;
    if (w->phase == 0) {
        fill_range(w->a, w->bytes, w->lane_begin, w->lane_end, GMO_PAT_SPLITMIX, w->seed);
    } else if (w->phase == 1) {
        verify_range(w->a, w->bytes, w->lane_begin, w->lane_end, GMO_PAT_SPLITMIX, w->seed, &w->res);
    } else {
        size_t off = (size_t)w->lane_begin * 16;
        size_t end = (size_t)w->lane_end * 16;
        if (end > w->bytes) end = w->bytes;
        memcpy(w->b + off, w->a + off, end - off);
    }
    return NULL;
}

static double wall(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int gmo_cpu_probe(size_t bytes, int threads, uint64_t seed, double seconds[3], gmo_verify_result* out) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    unsigned char* a = (unsigned char*)malloc(bytes ? bytes : 1);
    unsigned char* b = (unsigned char*)malloc(bytes ? bytes : 1);
    if (!a || !b) {
        free(a);
        free(b);
        return -1;
    }
    memset(a, 1, bytes); /* fault the pages in before timing */
    memset(b, 1, bytes);
    uint64_t lanes = (bytes + 15) / 16;
    pthread_t tid[256];
    worker_arg args[256];
    gmo_verify_result total = {0, UINT64_MAX, 0, 0};
    for (int phase = 0; phase < 3; ++phase) {
        double t0 = wall();
        for (int t = 0; t < threads; ++t) {
            args[t].a = a;
            args[t].b = b;
            args[t].bytes = bytes;
            args[t].lane_begin = lanes * (uint64_t)t / (uint64_t)threads;
            args[t].lane_end = lanes * (uint64_t)(t + 1) / (uint64_t)threads;
            args[t].seed = seed;
            args[t].phase = phase;
            pthread_create(&tid[t], NULL, worker, &args[t]);
        }
        for (int t = 0; t < threads; ++t) pthread_join(tid[t], NULL);
        seconds[phase] = wall() - t0;
        if (phase == 1) {
            for (int t = 0; t < threads; ++t) {
                total.mismatch_lanes += args[t].res.mismatch_lanes;
                if (args[t].res.first_bad_lane < total.first_bad_lane) total.first_bad_lane = args[t].res.first_bad_lane;
                total.xor64 ^= args[t].res.xor64;
                total.sum64 += args[t].res.sum64;
            }
        }
    }
    if (out) *out = total;
    free(a);
    free(b);
    return 0;
}
