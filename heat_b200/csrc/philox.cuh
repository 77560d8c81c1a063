// heat_b200/csrc/philox.cuh -- counter-based negative sampler (device + host twin).
//
// Replaces the reference's per-thread std::mt19937_64 sampler
// (negative_samplers/uniform_random_negative_sampler.cpp:10-36, random/uniform.hpp:16-30,
// seeded (epoch+1)*thread_id at train/engine.cpp:302) by Philox4x32-10 keyed on the run seed
// and counted by (global sample index, epoch, slot/4): thread-count and GPU-count
// independent, no sampler state in memory.
//
//   ctr  = { lo32(sample), hi32(sample), epoch, k >> 2 },  key = { lo32(seed), hi32(seed) }
//   draw = (u64(philox(ctr,key)[k & 3]) * num_items) >> 32
//
// The reference leaves slot k unchanged when its draw equals the positive item
// (uniform_random_negative_sampler.cpp:31-34; the vector is zero-initialised per epoch,
// engine.cpp:298).  Stateless form: walk back to the previous local sample until a draw
// differs from that sample's positive; before local sample 0 the slot holds 0.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define HEAT_HD __host__ __device__ __forceinline__
#else
#define HEAT_HD inline
#endif

namespace heat {

struct Philox4 { uint32_t v[4]; };

HEAT_HD uint32_t mulhi32(uint32_t a, uint32_t b)
{
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}

HEAT_HD Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1)
{
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t h0 = mulhi32(M0, c0), l0 = M0 * c0;
        const uint32_t h1 = mulhi32(M1, c2), l1 = M1 * c2;
        c0 = h1 ^ c1 ^ k0; c1 = l1;
        c2 = h0 ^ c3 ^ k1; c3 = l0;
        k0 += W0; k1 += W1;
    }
    Philox4 o; o.v[0] = c0; o.v[1] = c1; o.v[2] = c2; o.v[3] = c3;
    return o;
}

// Hot-tile mode (heat_config.tile_sampler with neg_sampler 1; the reference's RandomTileNegativeSampler::sampling,
// random_tile_negative_sampler.cpp:22-45 / paper Alg. 1, in stateless form): the samples of one refresh window
// (global sample index / refresh_interval) draw their negatives from a tile of `tile_size` item ids instead of the
// whole catalogue, so a window keeps touching the same rows.  Slot j of tile T of an epoch is itself a Philox draw,
//   item(T, j) = (u64(philox({lo32(T), hi32(T), epoch, j >> 2}, key ^ TILE)[j & 3]) * num_items) >> 32,
// and a sample's raw 32-bit draw x picks the slot j = (u64(x) * tile_size) >> 32.  tile_size 0 = uniform over all items.
constexpr uint32_t PHILOX_TILE_KEY0 = 0x656C6974u, PHILOX_TILE_KEY1 = 0x54414548u;   // "tile", "HEAT"

HEAT_HD uint32_t philox_pick(const Philox4 &r, uint32_t k)
{
    return (k & 3) == 0 ? r.v[0] : (k & 3) == 1 ? r.v[1] : (k & 3) == 2 ? r.v[2] : r.v[3];
}
HEAT_HD uint32_t map_draw(uint32_t x, uint64_t seed, uint32_t epoch, uint64_t sample, uint32_t num_items, uint32_t tile_size,
                          uint64_t refresh_interval)
{
    if (tile_size == 0) return mulhi32(x, num_items);
    const uint32_t j = mulhi32(x, tile_size);
    const uint64_t tile = sample / refresh_interval;
    const Philox4 r = philox4x32_10((uint32_t)tile, (uint32_t)(tile >> 32), epoch, j >> 2,
                                    (uint32_t)seed ^ PHILOX_TILE_KEY0, (uint32_t)(seed >> 32) ^ PHILOX_TILE_KEY1);
    return mulhi32(philox_pick(r, j), num_items);
}

HEAT_HD uint32_t philox_draw(uint64_t seed, uint32_t epoch, uint64_t sample, uint32_t k, uint32_t num_items,
                             uint32_t tile_size = 0, uint64_t refresh_interval = 1)
{
    const Philox4 r = philox4x32_10((uint32_t)sample, (uint32_t)(sample >> 32), epoch, k >> 2,
                                    (uint32_t)seed, (uint32_t)(seed >> 32));
    return map_draw(philox_pick(r, k), seed, epoch, sample, num_items, tile_size, refresh_interval);
}

}  // namespace heat
