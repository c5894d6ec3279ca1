// Height-field generation on the device: the sub-terrain generators the reference's
// Terrain class calls (utils/terrain.py:109-187: pyramid slope with or without uniform noise, pyramid stairs, discrete
// obstacles, stepping stones, gap, pit) and add_terrain_to_map (:147-164), one thread per height-field cell.
//
// The host keeps what must stay on the host: the np.random stream.  It draws every random number of the grid in the
// reference's order and hands the RESOLVED draws over (coarse noise levels, rectangle corners / sizes / heights, stone
// heights), together with a recipe per sub-terrain; the kernel turns recipes into cells.  Float arithmetic is the
// generators' own float64 arithmetic, operation for operation without contraction, so the field is bit-identical to
// the host generator's (tests/test_gpu_parity.py::test_device_terrain_generation_matches_the_host_generator).
#include "lgk_common.cuh"

__device__ __forceinline__ int16_t wrap16(int v) { return (int16_t)v; }   // int16 arithmetic of the numpy arrays

__device__ int terrain_cell(const LgkTerrainGrid& g, const LgkSubTerrain& t, const int32_t* __restrict__ draws,
                            const double* __restrict__ lin, int x, int y) {
  const int W = g.gen_w, L = g.gen_l;
  const int32_t* d = draws + t.draws_off;
  switch (t.type) {
    case LGK_TER_SLOPE:
    case LGK_TER_SLOPE_NOISE: {
      // pyramid_sloped_terrain: int16(peak * xx * yy), clipped to the platform height
      const int cx = W / 2, cy = L / 2;
      const double xx = __ddiv_rn((double)(cx - abs(cx - x)), (double)cx);
      const double yy = __ddiv_rn((double)(cy - abs(cy - y)), (double)cy);
      int h = (int)wrap16((int)__dmul_rn(__dmul_rn((double)t.p[0], xx), yy));   // astype(int16): toward zero
      h = min(max(h, t.p[1]), t.p[2]);
      if (t.type == LGK_TER_SLOPE_NOISE) {
        // random_uniform_terrain: bilinear up-sampling of the coarse levels, rounded half to even
        const int nx = t.p[3], ny = t.p[4];
        const double gx = lin[x], gy = lin[W + y];
        int x0 = nx > 1 ? min(max((int)floor(gx), 0), nx - 2) : 0;
        int y0 = ny > 1 ? min(max((int)floor(gy), 0), ny - 2) : 0;
        const double fx = __dsub_rn(gx, (double)x0), fy = __dsub_rn(gy, (double)y0);
        const int x1 = min(x0 + 1, nx - 1), y1 = min(y0 + 1, ny - 1);
        const double c00 = (double)d[x0 * ny + y0], c10 = (double)d[x1 * ny + y0];
        const double c01 = (double)d[x0 * ny + y1], c11 = (double)d[x1 * ny + y1];
        const double ox = __dsub_rn(1.0, fx), oy = __dsub_rn(1.0, fy);
        double f = __dmul_rn(__dmul_rn(c00, ox), oy);
        f = __dadd_rn(f, __dmul_rn(__dmul_rn(c10, fx), oy));
        f = __dadd_rn(f, __dmul_rn(__dmul_rn(c01, ox), fy));
        f = __dadd_rn(f, __dmul_rn(__dmul_rn(c11, fx), fy));
        h = (int)wrap16(h + (int)wrap16((int)rint(f)));
      }
      return h;
    }
    case LGK_TER_STAIRS: {
      // pyramid_stairs_terrain: ring k covers [k*sw, W - k*sw) x [k*sw, L - k*sw), k = 1..K
      const int sw = t.p[0];
      const int r = min(min(x / sw, (W - 1 - x) / sw), min(y / sw, (L - 1 - y) / sw));
      return (int)wrap16(t.p[1] * min(r, t.p[2]));
    }
    case LGK_TER_OBSTACLES: {
      // discrete_obstacles_terrain: rectangles in draw order, later ones on top, then the flat platform
      if (x >= t.p[1] && x < t.p[2] && y >= t.p[3] && y < t.p[4]) return 0;
      int h = 0;
      for (int k = 0; k < t.p[0]; ++k) {
        const int32_t* r = d + 5 * k;   // i, j, w, l, height
        if (x >= r[0] && x < r[0] + r[2] && y >= r[1] && y < r[1] + r[3]) h = r[4];
      }
      return h;
    }
    case LGK_TER_STONES: {
      // stepping_stones_terrain: stones of ss cells every ss + sd, everything else at the pit depth
      if (x >= t.p[4] && x < t.p[5] && y >= t.p[6] && y < t.p[7]) return 0;
      const int ss = t.p[0], pitch = t.p[0] + t.p[1];
      const int sa = x / pitch, sb = y / pitch;
      if (x - sa * pitch < ss && y - sb * pitch < ss) return d[sa * t.p[3] + sb];
      return t.p[2];
    }
    case LGK_TER_GAP: {
      const int cx = L / 2, cy = W / 2;   // (terrain.length // 2, terrain.width // 2, utils/terrain.py:169-170)
      if (x >= cx - t.p[2] && x < cx + t.p[2] && y >= cy - t.p[3] && y < cy + t.p[3]) return 0;
      if (x >= cx - t.p[0] && x < cx + t.p[0] && y >= cy - t.p[1] && y < cy + t.p[1]) return -1000;
      return 0;
    }
    case LGK_TER_PIT: {
      const int cx = L / 2, cy = W / 2;
      if (x >= cx - t.p[1] && x < cx + t.p[1] && y >= cy - t.p[1] && y < cy + t.p[1]) return -t.p[0];
      return 0;
    }
    default:
      return 0;
  }
}

__global__ void lgk_terrain_kernel(const __grid_constant__ LgkTerrainGrid g, const LgkSubTerrain* __restrict__ subs,
                                   const int32_t* __restrict__ draws, const double* __restrict__ lin,
                                   int16_t* __restrict__ field) {
  const int s = blockIdx.y;
  const int cell = blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= g.sub_len * g.sub_wid) return;
  const LgkSubTerrain t = subs[s];
  const int x = cell / g.sub_wid, y = cell - x * g.sub_wid;
  const int gx = g.border + t.row * g.sub_len + x, gy = g.border + t.col * g.sub_wid + y;   // add_terrain_to_map (:147-151)
  field[(size_t)gx * g.tot_cols + gy] = (int16_t)terrain_cell(g, t, draws, lin, x, y);
}

// env_origins z: max height of the 2 m x 2 m patch in the middle of the sub-terrain (:158-163)
__global__ void lgk_terrain_origin_kernel(const __grid_constant__ LgkTerrainGrid g, const LgkSubTerrain* __restrict__ subs,
                                          const int16_t* __restrict__ field, double* __restrict__ origin_z) {
  __shared__ int s_max[32];
  const LgkSubTerrain t = subs[blockIdx.x];
  const int pw = g.patch_y2 - g.patch_y1, n = (g.patch_x2 - g.patch_x1) * pw;
  int m = -32768;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int x = g.patch_x1 + i / pw, y = g.patch_y1 + i % pw;
    m = max(m, (int)field[(size_t)(g.border + t.row * g.sub_len + x) * g.tot_cols + g.border + t.col * g.sub_wid + y]);
  }
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(LGK_FULL_MASK, m, o));
  if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) m = max(m, s_max[w]);
    origin_z[t.row * g.num_cols + t.col] = __dmul_rn((double)m, g.vertical_scale);
  }
}

extern "C" int lgk_generate_terrain_impl(const LgkTerrainGrid* g, const LgkSubTerrain* subs, int n_sub, const int32_t* draws,
                                         const double* lin, int16_t* field, double* origin_z, cudaStream_t st) {
  LGK_CUDA_RETURN_IF(cudaMemsetAsync(field, 0, sizeof(int16_t) * (size_t)g->tot_rows * g->tot_cols, st));   // the border
  const int cells = g->sub_len * g->sub_wid;
  lgk_terrain_kernel<<<dim3((cells + 255) / 256, n_sub), 256, 0, st>>>(*g, subs, draws, lin, field);
  LGK_CUDA_RETURN_IF(cudaGetLastError());
  lgk_terrain_origin_kernel<<<n_sub, 256, 0, st>>>(*g, subs, field, origin_z);
  return (int)cudaGetLastError();
}
