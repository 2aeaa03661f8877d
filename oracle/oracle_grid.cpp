// TEST INFRASTRUCTURE ONLY -- CPU oracle: macrocell ranges / majorants and the synthetic noise volume.
#include "oracle_internal.h"

using namespace ovr_oracle;

extern "C" void ovr_oracle_macrocell_dims(const int32_t dims[3], int32_t mc_dims[3])
{
  // sp_singlemc.cu:99-111: mcDims = ceil(dims / 16)
  for (int i = 0; i < 3; i++) mc_dims[i] = (dims[i] + MC_SIZE - 1) / MC_SIZE;
}

extern "C" void ovr_oracle_build_ranges(const float* voxels, const int32_t dims[3], int mode, float* ranges)
{
  int32_t mc[3];
  ovr_oracle_macrocell_dims(dims, mc);
  const size_t sy = size_t(dims[0]), sz = size_t(dims[0]) * size_t(dims[1]);
  parallel_rows(0, mc[2], 1, 0, [&](int cz) {
    for (int cy = 0; cy < mc[1]; cy++)
      for (int cx = 0; cx < mc[0]; cx++) {
        const int c[3] = { cx, cy, cz };
        int b[3], e[3]; // voxel range [b, e)
        for (int a = 0; a < 3; a++) {
          if (mode == OVR_ORACLE_MC_REFERENCE) {
            // sp_singlemc.cu:36-42: begin = max(16c-1, 0); end = min(begin + 16 + 1, dims)
            b[a] = std::max(c[a] * MC_SIZE - 1, 0);
            e[a] = std::min(b[a] + MC_SIZE + 1, dims[a]);
          }
          else {
            // what the B200 device builds: voxels 16c-2 .. 16c+17 (tri-lerp footprint of every position in the
            // cell plus half a voxel of slack below and 1.5 above for fp error in the DDA)
            b[a] = std::max(c[a] * MC_SIZE - 2, 0);
            e[a] = std::min(c[a] * MC_SIZE + MC_SIZE + 2, dims[a]);
          }
        }
        float lo = INFINITY, hi = -INFINITY; // range1f default = empty interval
        for (int z = b[2]; z < e[2]; z++)
          for (int y = b[1]; y < e[1]; y++) {
            const float* row = voxels + size_t(y) * sy + size_t(z) * sz;
            for (int x = b[0]; x < e[0]; x++) {
              lo = std::min(lo, row[x]);
              hi = std::max(hi, row[x]);
            }
          }
        const size_t idx = size_t(cx) + size_t(mc[0]) * (size_t(cy) + size_t(mc[1]) * size_t(cz));
        ranges[2 * idx] = lo;
        ranges[2 * idx + 1] = hi;
      }
  });
}

extern "C" void ovr_oracle_build_majorants(const float* ranges, int64_t n_cells, const float* tf_alpha, int n_alpha,
                                           float win_lo, float win_hi, float* majorants)
{
  // sp_singlemc.cu:56-97 with the three defects of SURVEY A.7-q7 fixed: the index is clamped before the
  // float->uint conversion, every thread sees the whole table, and the window is the lookup window.
  const float rcp = 1.f / (win_hi - win_lo);
  for (int64_t i = 0; i < n_cells; i++) {
    const float lo = (cclamp(ranges[2 * i], win_lo, win_hi) - win_lo) * rcp;
    const float hi = (cclamp(ranges[2 * i + 1], win_lo, win_hi) - win_lo) * rcp;
    float fl = std::floor(std::fmaf(lo, float(n_alpha - 1), 0.5f)) - 1.f;
    float fh = std::floor(std::fmaf(hi, float(n_alpha - 1), 0.5f)) + 1.f;
    fl = cclamp(fl, 0.f, float(n_alpha - 1));
    fh = cclamp(fh, 0.f, float(n_alpha - 1));
    float m = 0.f;
    for (int k = int(fl); k <= int(fh); k++) m = std::max(m, tf_alpha[k]);
    majorants[i] = m;
  }
}

// ---- synthetic volume noise(N, seed): SURVEY 8(d). Integer-exact so that CPU and GPU agree bit for bit. ----
static inline uint32_t fmix32(uint32_t h)
{
  h ^= h >> 16;
  h *= 0x85ebca6bu;
  h ^= h >> 13;
  h *= 0xc2b2ae35u;
  h ^= h >> 16;
  return h;
}
static inline uint64_t lattice(uint32_t seed, uint32_t a, uint32_t b, uint32_t c, uint32_t o)
{
  return fmix32(seed ^ (a * 73856093u) ^ (b * 19349663u) ^ (c * 83492791u) ^ (o * 0x9E3779B9u)) >> 16;
}
static inline float noise_voxel(uint32_t seed, uint32_t x, uint32_t y, uint32_t z)
{
  uint64_t sum = 0; // sum_o (8>>o) * O_o * (2^18 / P^3), O_o = integer trilinear blend of 16-bit lattice values
  for (uint32_t o = 0; o < 4; o++) {
    const uint32_t sh = 6 - o, P = 1u << sh;
    const uint32_t a = x >> sh, b = y >> sh, c = z >> sh;
    const uint64_t fx = x & (P - 1), fy = y & (P - 1), fz = z & (P - 1);
    uint64_t acc = 0;
    for (uint32_t k = 0; k < 8; k++) {
      const uint32_t dx = k & 1, dy = (k >> 1) & 1, dz = k >> 2;
      const uint64_t w = (dx ? fx : P - fx) * (dy ? fy : P - fy) * (dz ? fz : P - fz);
      acc += w * lattice(seed, a + dx, b + dy, c + dz, o);
    }
    sum += uint64_t(8u >> o) * (acc << (3 * o));
  }
  return float(double(sum) / double(15ull * 65536ull * 262144ull));
}

extern "C" void ovr_oracle_noise(int32_t n, uint32_t seed, const int32_t org[3], const int32_t bdims[3], float* out,
                                 int num_threads)
{
  (void)n;
  parallel_rows(0, bdims[2], 1, num_threads, [&](int z) {
    for (int y = 0; y < bdims[1]; y++) {
      float* row = out + (size_t(z) * size_t(bdims[1]) + size_t(y)) * size_t(bdims[0]);
      for (int x = 0; x < bdims[0]; x++)
        row[x] = noise_voxel(seed, uint32_t(org[0] + x), uint32_t(org[1] + y), uint32_t(org[2] + z));
    }
  });
}

// Sparse (foveated) sampling mask: generate_blue_noise_kernel (ovr/common/random/blue_noise.h:81-101) followed by
// generate_sparse_samples (ovr/common/generate_mask.cu:55-86). keep[i] = 1 when pixel i = x + y * width is rendered;
// p_out (may be NULL) receives the acceptance probability, margin information for the tests: the reference evaluates
// exp() with the device's approximate __expf and contracted multiply-adds, so a pixel whose noise value lies within a few
// ulp of p can legitimately fall on either side on the GPU.
extern "C" void ovr_oracle_sparse_mask(int width, int height, int frame_index, const float* noise, int tile_xy, int tile_t,
                                       const float mean[2], float sigma, float base_noise, uint8_t* keep, float* p_out)
{
  const float sigma_rcp2 = 1.f / (sigma * sigma); // generate_mask.cu:92
  const float aspect = (float)width / (float)height;
  for (int y = 0; y < height; y++)
    for (int x = 0; x < width; x++) {
      const size_t i = (size_t)x + (size_t)y * (size_t)width;
      const float d = noise[((size_t)(y % tile_xy) * tile_xy + (size_t)(x % tile_xy)) * tile_t + (size_t)(frame_index % tile_t)];
      const float fx = (float)x / (float)width - mean[0];
      const float fy = ((float)y / (float)height - mean[1]) / aspect;
      const float p = (1.0f - base_noise) * std::exp(-0.5f * (fx * fx + fy * fy) * sigma_rcp2) + base_noise;
      keep[i] = d < p ? 1 : 0;
      if (p_out) p_out[i] = p;
    }
}
