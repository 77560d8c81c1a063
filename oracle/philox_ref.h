/* oracle/philox_ref.h -- TEST INFRASTRUCTURE (CPU oracle), not product code.
 *
 * Plain-C restatement of the negative-sampling contract of the B200 path.
 * The reference draws negatives with std::mt19937_64 seeded per OpenMP thread
 * (uniform_random_negative_sampler.cpp:10-36, uniform.hpp:16-30, seed at
 * engine.cpp:302), which is neither thread-count independent nor reproducible
 * on a GPU.  The north-star replaces it by a counter-based Philox4x32-10
 * stream that this header defines for the host side; the CUDA kernel carries
 * an independent device implementation and tests/ compare the two bit for bit.
 *
 * Contract ("HEAT-B200 sampler v1"):
 *   ctr  = { lo32(sample), hi32(sample), (u32)epoch, (u32)(k >> 2) }
 *   key  = { lo32(seed), hi32(seed) }
 *   r    = philox4x32_10(ctr, key)[k & 3]
 *   draw = (u64(r) * num_items) >> 32                 -- in [0, num_items)
 * and the reference's "ignore_pos" quirk (slot keeps its previous value when
 * the draw equals the positive, uniform_random_negative_sampler.cpp:26-36,
 * vector zero-initialised per epoch at engine.cpp:298) becomes the stateless
 * look-back rule implemented by heat_oracle_neg_id() below.
 */
#ifndef HEAT_ORACLE_PHILOX_REF_H
#define HEAT_ORACLE_PHILOX_REF_H
#include <stdint.h>

static inline void heat_oracle_philox4x32_10(const uint32_t ctr_in[4],
                                             const uint32_t key_in[2],
                                             uint32_t out[4])
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    const uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    uint32_t c0 = ctr_in[0], c1 = ctr_in[1], c2 = ctr_in[2], c3 = ctr_in[3];
    uint32_t k0 = key_in[0], k1 = key_in[1];
    for (int round = 0; round < 10; ++round) {
        uint64_t p0 = (uint64_t)M0 * c0;
        uint64_t p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* One raw draw in [0, num_items). */
static inline uint64_t heat_oracle_draw(uint64_t seed, uint64_t epoch,
                                        uint64_t sample, uint32_t k,
                                        uint64_t num_items)
{
    uint32_t ctr[4] = { (uint32_t)sample, (uint32_t)(sample >> 32),
                        (uint32_t)epoch, k >> 2 };
    uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
    uint32_t r[4];
    heat_oracle_philox4x32_10(ctr, key, r);
    return ((uint64_t)r[k & 3] * num_items) >> 32;
}

/* Hot-tile mode (heat_config.tile_sampler; the reference's RandomTileNegativeSampler::sampling,
 * random_tile_negative_sampler.cpp:22-45, in stateless form): the samples of one window of
 * refresh_interval global sample indices draw from a tile of tile_size item ids.
 *   j    = (u64(r) * tile_size) >> 32                         -- r = the sample's raw 32-bit draw
 *   T    = sample / refresh_interval
 *   t    = philox4x32_10({lo32(T), hi32(T), epoch, j >> 2}, {lo32(seed) ^ "tile", hi32(seed) ^ "HEAT"})[j & 3]
 *   draw = (u64(t) * num_items) >> 32
 * tile_size 0 = the uniform contract above. */
static inline uint64_t heat_oracle_draw_tile(uint64_t seed, uint64_t epoch, uint64_t sample, uint32_t k,
                                             uint64_t num_items, uint64_t tile_size, uint64_t refresh_interval)
{
    uint32_t ctr[4] = { (uint32_t)sample, (uint32_t)(sample >> 32), (uint32_t)epoch, k >> 2 };
    uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
    uint32_t r[4];
    heat_oracle_philox4x32_10(ctr, key, r);
    if (tile_size == 0) return ((uint64_t)r[k & 3] * num_items) >> 32;
    const uint32_t j = (uint32_t)(((uint64_t)r[k & 3] * tile_size) >> 32);
    const uint64_t T = sample / (refresh_interval ? refresh_interval : 1);
    uint32_t ctr2[4] = { (uint32_t)T, (uint32_t)(T >> 32), (uint32_t)epoch, j >> 2 };
    uint32_t key2[2] = { (uint32_t)seed ^ 0x656C6974u, (uint32_t)(seed >> 32) ^ 0x54414548u };
    heat_oracle_philox4x32_10(ctr2, key2, r);
    return ((uint64_t)r[j & 3] * num_items) >> 32;
}
static inline uint64_t heat_oracle_neg_id_tile(uint64_t seed, uint64_t epoch, uint64_t sample_offset, const uint64_t *clicks,
                                               uint64_t s, uint32_t k, uint64_t num_items, uint64_t tile_size,
                                               uint64_t refresh_interval)
{
    for (;;) {
        uint64_t d = heat_oracle_draw_tile(seed, epoch, sample_offset + s, k, num_items, tile_size, refresh_interval);
        if (d != clicks[2 * s + 1]) return d;
        if (s == 0) return 0;
        --s;
    }
}

/* Negative id of slot k for local sample s (global index sample_offset + s).
 * clicks is the [N,2] u64 (user, item) array of the local shard. */
static inline uint64_t heat_oracle_neg_id(uint64_t seed, uint64_t epoch,
                                          uint64_t sample_offset,
                                          const uint64_t *clicks, uint64_t s,
                                          uint32_t k, uint64_t num_items)
{
    for (;;) {
        uint64_t d = heat_oracle_draw(seed, epoch, sample_offset + s, k, num_items);
        if (d != clicks[2 * s + 1]) return d;
        if (s == 0) return 0;   /* fresh zero-initialised vector, engine.cpp:298 */
        --s;
    }
}
#endif
