// ORACLE (test infrastructure only — never linked into the product library).
//
// Philox4x32-10 counter-based RNG (Salmon et al., SC'11, Random123 v1.x).
// The reference (MonteCarlo.jl) draws from Julia's global MersenneTwister
// (src/flavors/DQMC/DQMC.jl:573, src/flavors/MC/MC.jl:324); that stream cannot
// be reproduced without Julia, so SURVEY.md §8(d) / Appendix D define the
// stream both the oracle and the CUDA library must consume:
//   DQMC : key = (seed, global chain id), counter = (sweep, visit, site, 0)
//          u = ((x1:x0) >> 11) * 2^-53   in [0,1)  (53-bit, like Julia rand())
//   Ising: key = (seed, global replica id), counter = (sweep, colour, group, 0)
// Pinned against the Random123 known-answer vectors in tests/test_philox.py.
#pragma once
#include <stdint.h>

static inline void philox4x32_10_ref(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static inline double philox_uniform53_ref(uint32_t lo, uint32_t hi)
{
    uint64_t x = ((uint64_t)hi << 32) | lo;
    return (double)(x >> 11) * (1.0 / 9007199254740992.0);
}

// DQMC acceptance uniform for (sweep, visit, site) of chain `chain`.
static inline double dqmc_uniform_ref(uint64_t seed, uint32_t chain, uint32_t sweep,
                                      uint32_t visit, uint32_t site)
{
    uint32_t ctr[4] = {sweep, visit, site, 0u};
    uint32_t key[2] = {(uint32_t)seed ^ (uint32_t)(seed >> 32), chain};
    uint32_t o[4];
    philox4x32_10_ref(ctr, key, o);
    return philox_uniform53_ref(o[0], o[1]);
}
