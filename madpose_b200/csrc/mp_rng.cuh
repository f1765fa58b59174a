// std::mt19937 + the libstdc++ distributions the reference's control loop draws from,
// reproduced on the device so that the GPU loop consumes exactly the random streams of
// the CPU loop: stream A = HybridUniformSampling::rng_
// (ext/RansacLib/RansacLib/hybrid_sampling.h:52,134), stream B = the loop's own engine
// (hybrid_ransac.h:79-80: solver selection and every LO shuffle).
// The generator state IS a ring of raw (untempered) words: word r >= 624 is
// x[r] = x[r-227] ^ twist(x[r-624], x[r-623]); keeping the last MT_RING raw words lets the
// loop rewind a stream by up to MT_RING-624 outputs (speculative batches of iterations).
#pragma once
#include "mp_common.cuh"

namespace mp {

constexpr int MT_N = 624, MT_M = 397, MT_RING = 10 * MT_N; // >= 20 words x LP_BMAX iterations of a batch + 624

struct alignas(128) MtStream { // line-aligned: rings of neighbouring pairs never share a cache line
    uint32_t x[MT_RING];
    unsigned long long gen; // raw words generated so far (multiple of 624, >= 624)
    unsigned long long pos; // raw index of the next output word (>= 624)
};

// seeds raw words 0..623 (std::mersenne_twister_engine::seed)
MP_HD void mt_seed(MtStream *s, uint32_t seed) {
    s->x[0] = seed;
    for (int i = 1; i < MT_N; i++) s->x[i] = 1812433253u * (s->x[i - 1] ^ (s->x[i - 1] >> 30)) + (uint32_t)i;
    s->gen = MT_N; s->pos = MT_N;
}
MP_HD uint32_t mt_twist(uint32_t a, uint32_t b, uint32_t c) {
    const uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
    return c ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
}
MP_HD uint32_t mt_temper(uint32_t y) {
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    return y;
}
// generates raw words gen .. gen+623.  `lane`/`nlanes` spread each of the three
// dependency-free phases over cooperating threads; SYNC separates the phases
// (a no-op functor for a single thread).
template <class Sync> MP_HD void mt_refill(MtStream *s, int lane, int nlanes, Sync sync) {
    const unsigned long long g = s->gen;
    const unsigned gb = (unsigned)(g % MT_RING); // ring slot of raw word g; all other slots by 32-bit add + one wrap
    auto slot = [gb](int off) { // ring slot of raw word g + off, -MT_N <= off < MT_N
        unsigned r = gb + (unsigned)(off + MT_RING);
        if (r >= 2u * MT_RING) r -= 2u * MT_RING;
        else if (r >= (unsigned)MT_RING) r -= MT_RING;
        return r;
    };
    const int bounds[4] = {0, MT_N - MT_M, 2 * (MT_N - MT_M), MT_N};
    for (int ph = 0; ph < 3; ph++) {
        for (int i = bounds[ph] + lane; i < bounds[ph + 1]; i += nlanes)
            s->x[slot(i)] = mt_twist(s->x[slot(i - MT_N)], s->x[slot(i - MT_N + 1)], s->x[slot(i - MT_N + MT_M)]);
        sync();
    }
    if (lane == 0) s->gen = g + MT_N;
    sync();
}
struct NoSync { MP_HD void operator()() const {} };

// single-thread draw
MP_HD uint32_t mt_next(MtStream *s) {
    if (s->pos == s->gen) mt_refill(s, 0, 1, NoSync());
    const uint32_t y = s->x[s->pos % MT_RING];
    s->pos++;
    return mt_temper(y);
}
// Word sources: anything with `uint32_t next()` returning the next TEMPERED output word.
struct MtDirect { // the stream itself (ring in global memory)
    MtStream *s;
    MP_HD uint32_t next() { return mt_next(s); }
    MP_HD unsigned long long position() const { return s->pos; }
};
// A window of raw words [pos0, pos0 + cnt) copied out of the ring (shared memory on the device):
// consumed sequentially by one thread; falls back to the stream when the window is used up.
struct MtWindow {
    MtStream *s;
    const uint32_t *w;
    unsigned long long pos0;
    int cnt, used;
    bool spilled;
    MP_HD uint32_t next() {
        if (!spilled) {
            if (used < cnt) return mt_temper(w[used++]);
            spilled = true;
            s->pos = pos0 + (unsigned long long)used;
        }
        return mt_next(s);
    }
    MP_HD unsigned long long position() const { return spilled ? s->pos : pos0 + (unsigned long long)used; }
};

// Read-only window for a thread that draws at a GUESSED offset next to other threads: never touches the stream; past
// the end it flags `overflow` and hands out changing words so that the retry loops of the caller still end.
struct MtWindowRO {
    const uint32_t *w;
    int cnt, used;
    bool overflow;
    MP_HD uint32_t next() {
        if (used < cnt) return mt_temper(w[used++]);
        overflow = true;
        return 2654435761u * (uint32_t)(++used);
    }
};

// std::uniform_int_distribution<int>(a, b) on a 32-bit engine (libstdc++ 13: Lemire's method)
template <class Src> MP_HD int mt_uniform_int_t(Src &s, int a, int b) {
    const uint32_t range = (uint32_t)(b - a) + 1u; // never 0 here (b - a < 2^32 - 1)
    unsigned long long prod = (unsigned long long)s.next() * range;
    uint32_t low = (uint32_t)prod;
    if (low < range) {
        const uint32_t thr = (0u - range) % range;
        while (low < thr) {
            prod = (unsigned long long)s.next() * range;
            low = (uint32_t)prod;
        }
    }
    return a + (int)(prod >> 32);
}
MP_HD int mt_uniform_int(MtStream *s, int a, int b) {
    MtDirect d{s};
    return mt_uniform_int_t(d, a, b);
}
// would the word `w` be rejected by uniform_int with this range?
MP_HD bool mt_lemire_rejects(uint32_t w, uint32_t range) {
    const uint32_t low = (uint32_t)((unsigned long long)w * range);
    return low < range && low < (0u - range) % range;
}
// std::uniform_real_distribution<double>(0, hi): generate_canonical<double, 53> = two words
template <class Src> MP_HD double mt_uniform_real_t(Src &s, double hi) {
    const double lo32 = (double)s.next(), hi32 = (double)s.next();
    double c = (lo32 + hi32 * 4294967296.0) / 18446744073709551616.0;
    if (c >= 1.0) c = 1.0 - 1.1102230246251565e-16; // nextafter(1, 0)
    return c * hi; // (b - a) * c + a with a = 0
}
MP_HD double mt_uniform_real(MtStream *s, double hi) {
    MtDirect d{s};
    return mt_uniform_real_t(d, hi);
}

// HybridUniformSampling::DrawSample / ShuffleSample for one data type (hybrid_sampling.h:84-131):
// k distinct indices of [0, n).  scratch: n ints, only touched on the Fisher-Yates branch.
template <class Src> MP_HD void mt_draw_sample_t(Src &s, int n, int k, int *out, int *scratch) {
    if (k <= 0) return;
    const double coeff = (double)n / (double)(n - k);
    if (coeff < 2.718281828459045235360287471352662498) {
        for (int i = 0; i < k; i++) {
            bool dup = true;
            while (dup) {
                dup = false;
                out[i] = mt_uniform_int_t(s, 0, n - 1);
                for (int j = 0; j < i; j++)
                    if (out[j] == out[i]) { dup = true; break; }
            }
        }
    } else {
        for (int i = 0; i < n; i++) scratch[i] = i;
        if (k != n)
            for (int i = 0; i < n - 1; i++) {
                const int j = mt_uniform_int_t(s, i, n - 1);
                const int t = scratch[i]; scratch[i] = scratch[j]; scratch[j] = t;
            }
        for (int i = 0; i < k; i++) out[i] = scratch[i];
    }
}
// words mt_draw_sample_t takes when no draw has to be repeated
MP_HD int mt_draw_words(int n, int k) {
    if (k <= 0) return 0;
    const double coeff = (double)n / (double)(n - k);
    if (coeff < 2.718281828459045235360287471352662498) return k;
    return k != n ? n - 1 : 0;
}
MP_HD void mt_draw_sample(MtStream *s, int n, int k, int *out, int *scratch) {
    MtDirect d{s};
    mt_draw_sample_t(d, n, k, out, scratch);
}

} // namespace mp
