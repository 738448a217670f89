// Host stand-in for the cuRAND device API used by the reference (curand_init, curand_uniform).
// Two modes per state: (a) replay of a caller-supplied uniform sequence (known-answer tests),
// (b) a per-state xoshiro128++ stream seeded like the reference seeds XORWOW: by pixel index.
// curand_uniform returns values in (0,1] exactly like cuRAND's contract.
#pragma once
#include <cstdint>
struct curandState {
    const float *replay = nullptr; int replay_n = 0; int replay_pos = 0;
    uint32_t s[4] = {1, 2, 3, 4};
    long long draws = 0;
};
typedef curandState curandStateXORWOW_t;
typedef curandState curandStateXORWOW;

static inline uint32_t shim_rotl(uint32_t x, int k) { return (x << k) | (x >> (32 - k)); }
static inline uint64_t shim_splitmix(uint64_t &x) {
    uint64_t z = (x += 0x9e3779b97f4a7c15ULL);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL; z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
static inline void curand_init(unsigned long long seed, unsigned long long seq, unsigned long long off, curandState *st) {
    uint64_t x = seed * 0x2545F4914F6CDD1DULL + seq * 0x9E3779B97F4A7C15ULL + off + 0x1234567ULL;
    uint64_t a = shim_splitmix(x), b = shim_splitmix(x);
    st->s[0] = (uint32_t)a; st->s[1] = (uint32_t)(a >> 32); st->s[2] = (uint32_t)b; st->s[3] = (uint32_t)(b >> 32);
    if (!(st->s[0] | st->s[1] | st->s[2] | st->s[3])) st->s[0] = 1;
    st->replay = nullptr; st->replay_n = st->replay_pos = 0; st->draws = 0;
}
static inline uint32_t shim_next(curandState *st) {
    uint32_t *s = st->s;
    const uint32_t result = shim_rotl(s[0] + s[3], 7) + s[0];
    const uint32_t t = s[1] << 9;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = shim_rotl(s[3], 11);
    return result;
}
static inline float curand_uniform(curandState *st) {
    st->draws++;
    if (st->replay) { float v = st->replay[st->replay_pos % st->replay_n]; st->replay_pos++; return v; }
    return ((shim_next(st) >> 8) + 1u) * (1.0f / 16777216.0f);   // (0,1]
}
