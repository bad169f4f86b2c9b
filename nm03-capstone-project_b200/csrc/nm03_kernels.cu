// nm03_kernels.cu — sm_100a kernels of the per-slice brain-MRI pipeline.
//
//   K1  k1_preprocess   IntensityNormalization + IntensityClipping + VectorMedianFilter +
//                       ImageSharpening (+ the region-growing range test) in one pass over the raw
//                       int16 slice  (reference main_sequential.cpp:195-210, :233).
//   K23 k23_segment     SeededRegionGrowing + ImageCaster + Dilation/Erosion on bit-packed planes
//                       held in (distributed) shared memory, one thread-block cluster per slice
//                       (reference main_sequential.cpp:232-252, test_pipeline.cpp:119-125).
//
// Nothing here is a dense contraction, so no tensor cores: K1 is bound by the integer min/max
// pipe (packed VIMNMX.S16x2 networks), K23 by shared-memory latency.  Built with -fmad=false; the
// places where rounding must follow the oracle use explicit _rn intrinsics.
#include "nm03_k1.cuh"

namespace nm03 {

// f() for all 65536 keys, computed once per parameter set with exactly the operations above; K1 then
// replaces ~30 instructions per pixel (IEEE division with its special-case scaffolding) by one
// cached load.  lut index = the key's 16 bits as an unsigned number.
__global__ void k_build_lut(const __grid_constant__ PreParams p, float* __restrict__ lut) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 65536) lut[i] = key_to_float((int)(short)i, p);
}

cudaError_t launch_build_lut(const PreParams& p, float* lut, cudaStream_t s) {
  k_build_lut<<<256, 256, 0, s>>>(p, lut);
  return cudaGetLastError();
}

// K1 is instantiated per median window in its own translation unit (nm03_k1_w.cu with -DNM03_K1_WINDOW=k)
#define NM03_K1_WINDOWS(X) X(1) X(3) X(5) X(7) X(9) X(11) X(13) X(15)
#define NM03_DECL(k) cudaError_t launch_k1_w##k(const PreParams& p, int sharpen_half, int mode, cudaStream_t s); cudaError_t config_k1_w##k();
NM03_K1_WINDOWS(NM03_DECL)
#undef NM03_DECL

int median_network_ops(int median_window) {
  switch (median_window) {
    case 3: return NM03_MEDIAN_OPS_3;
    case 5: return NM03_MEDIAN_OPS_5;
    case 7: return NM03_MEDIAN_OPS_7;
    case 9: return NM03_MEDIAN_OPS_9;
    default: return 0;
  }
}

cudaError_t launch_preprocess(const PreParams& p, int median_window, int sharpen_half, int mode, cudaStream_t s) {
  switch (median_window) {
#define NM03_CASE(k) case k: return launch_k1_w##k(p, sharpen_half, mode, s);
    NM03_K1_WINDOWS(NM03_CASE)
#undef NM03_CASE
    default: return cudaErrorInvalidValue;
  }
}

// =================================================================================================
// Windows above 15x15: generic median of the raw samples, one output per thread
// =================================================================================================
// The tile halo of K1 ends at 15x15.  FAST's VectorMedianFilter takes any odd window (main_sequential.cpp:204), so larger
// ones go through two passes: this kernel writes the K x K median of the RAW 16-bit samples (coordinates clamped to the
// edge, nm03o_median) and K1 runs on that plane with a 1x1 window — normalise + clip is monotone, so it commutes with
// the median (the same identity K1 itself rests on).  Exact for every input, ties included, and deliberately simple:
// a bound on the window's values, then a most-significant-bit-first search for the sample of rank (K*K+1)/2, counting
// the samples below a trial value.  About 3 * (2 + bits that differ) * K*K instructions per pixel, read through L1.
__global__ void __launch_bounds__(256) k_median_big(const uint16_t* __restrict__ in, uint16_t* __restrict__ out, int w, int h, int R,
                                                    uint32_t flip) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= w || y >= h) return;
  const uint16_t* __restrict__ img = in + (size_t)blockIdx.z * w * h;
  const int K = 2 * R + 1;
  const uint32_t rank = (uint32_t)(K * K + 1) / 2;
  // unsigned order: signed samples are flipped by 0x8000
  uint32_t mn = 0xffffu, mx = 0u;
  for (int dy = -R; dy <= R; ++dy) {
    const uint16_t* __restrict__ row = img + (size_t)min(max(y + dy, 0), h - 1) * w;
    for (int dx = -R; dx <= R; ++dx) {
      const uint32_t v = (uint32_t)__ldg(row + min(max(x + dx, 0), w - 1)) ^ flip;
      mn = min(mn, v);
      mx = max(mx, v);
    }
  }
  uint32_t cand = mn;
  if (mn != mx) {
    const int hb = 31 - __clz((int)(mn ^ mx));   // every sample shares the bits above hb
    cand = mn & ~((2u << hb) - 1u);
    for (int b = hb; b >= 0; --b) {
      const uint32_t t = cand | (1u << b);
      uint32_t below = 0u;
      for (int dy = -R; dy <= R; ++dy) {
        const uint16_t* __restrict__ row = img + (size_t)min(max(y + dy, 0), h - 1) * w;
        for (int dx = -R; dx <= R; ++dx)
          below += (((uint32_t)__ldg(row + min(max(x + dx, 0), w - 1)) ^ flip) < t) ? 1u : 0u;
      }
      if (below < rank) cand = t;   // fewer than `rank` samples below t: the median is at least t
    }
  }
  out[(size_t)blockIdx.z * w * h + (size_t)y * w + x] = (uint16_t)(cand ^ flip);
}

cudaError_t launch_median_big(const void* in, int n, int w, int h, int window, int input_unsigned, void* out, cudaStream_t s) {
  if (window < 3 || (window & 1) == 0 || window > BIG_MAX_WINDOW || n <= 0) return cudaErrorInvalidValue;
  const int gy = (h + 7) / 8;
  if (gy > 65535) return cudaErrorInvalidValue;
  for (int z0 = 0; z0 < n; z0 += 65535) {   // gridDim.z limit
    const int nz = n - z0 < 65535 ? n - z0 : 65535;
    const dim3 grid((unsigned)((w + 31) / 32), (unsigned)gy, (unsigned)nz);
    const size_t off = (size_t)z0 * w * h;
    k_median_big<<<grid, 256, 0, s>>>(reinterpret_cast<const uint16_t*>(in) + off, reinterpret_cast<uint16_t*>(out) + off, w, h,
                                      window / 2, input_unsigned ? 0u : 0x8000u);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// =================================================================================================
// K23
// =================================================================================================
// Horizontal flood inside one 32-pixel word: spread the bits of c (c subset of p) along the runs of
// ones of p, in both directions.  Adding c to p makes the carry ripple to the end of each run.
__device__ __forceinline__ uint32_t hfill(uint32_t c, uint32_t p) {
  const uint32_t up = ((p + c) ^ p) & p;
  const uint32_t rp = __brev(p), rc = __brev(c);
  const uint32_t dn = __brev(((rp + rc) ^ rp) & rp);
  return up | dn | c;
}

__device__ __forceinline__ uint32_t spread4(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }

// write 32 mask pixels (one word) as uint8 {0,1}
__device__ __forceinline__ void store_word_u8(uint8_t* __restrict__ row, int x, int w, uint32_t bits, bool vec) {
  if (vec && x + 32 <= w) {
    uint4 a, b;
    a.x = spread4(bits & 15u); a.y = spread4((bits >> 4) & 15u); a.z = spread4((bits >> 8) & 15u); a.w = spread4((bits >> 12) & 15u);
    b.x = spread4((bits >> 16) & 15u); b.y = spread4((bits >> 20) & 15u); b.z = spread4((bits >> 24) & 15u); b.w = spread4(bits >> 28);
    *reinterpret_cast<uint4*>(row + x) = a;
    *reinterpret_cast<uint4*>(row + x + 16) = b;
  } else {
    for (int i = 0; i < 32 && x + i < w; ++i) row[x + i] = (bits >> i) & 1u;
  }
}

// One sweep pass of the flood fill over the rows this CTA owns.  A thread owns one 32-pixel word
// column over a segment of rows and walks it down, then up; the freshly computed word of the previous
// row travels in a register, the other vertical neighbour and the two horizontal neighbour words are
// read from shared memory as they are (chaotic relaxation of a monotone operator: any interleaving
// converges to the same fixed point, and a pass without any change proves it has been reached).
// MASKED = image-border pixels may be labelled but never act as sources (rg_border_expand = 0).
// Rows of both bit planes are stored with an ODD stride in shared memory, so a warp whose lanes walk
// along the words of a row and a warp whose lanes walk down the rows of one word column are both
// free of bank conflicts.
__host__ __device__ inline int seg_row_stride(int wpr) { return wpr | 1; }

// Activity flags: an item (row segment x word column) is processed in a pass only if it, or one of
// the items it reads from, changed since it was last looked at; a changing item wakes itself and its
// neighbours for the next pass.  Work is therefore proportional to the grown region, not to the image.
template <bool CONN8, bool MASKED>
__device__ __forceinline__ int fill_pass(uint32_t* __restrict__ Mm, const uint32_t* __restrict__ Pp,
                                         const SegParams& p, int nrows, int y_begin, int tid,
                                         uint8_t* __restrict__ act_now, uint8_t* __restrict__ act_next) {
  const int wpr = p.wpr, ws = seg_row_stride(p.wpr);
#define MROW(r) (Mm + (size_t)((r) + SEG_HALO) * ws)
  const int lastk = (p.w - 1) >> 5;
  const uint32_t lastbit = 1u << ((p.w - 1) & 31);
  int changed = 0;
  const int nseg = (nrows + p.seg_rows - 1) / p.seg_rows;
  for (int it = tid; it < nseg * wpr; it += SEG_THREADS) {
    if (!act_now[it]) continue;
    act_now[it] = 0;
    int item_changed = 0;
    const int sgi = it / wpr, k = it - sgi * wpr;
    const int ra = sgi * p.seg_rows, rb = min(nrows, ra + p.seg_rows);
    const bool has_l = k > 0, has_r = k < wpr - 1;
    uint32_t colmask = 0xffffffffu, lmask = 0xffffffffu, rmask = 0xffffffffu;  // source masks: own, left, right word
    if (MASKED) {
      if (k == 0) colmask &= ~1u;
      if (k == lastk) colmask &= ~lastbit;
      if (k - 1 == 0) lmask &= ~1u;
      if (k - 1 == lastk) lmask &= ~lastbit;
      if (k + 1 == lastk) rmask &= ~lastbit;
    }
    auto row_is_source = [&](int rr) { return !MASKED || (y_begin + rr > 0 && y_begin + rr < p.h - 1); };
    // words k-1, k, k+1 of one mask row: c = raw centre word, s/l/r = the three words as SOURCES
    struct Row3 { uint32_t c, s, l, r; };
    auto load_row = [&](int rr) {
      const uint32_t* row = MROW(rr);
      Row3 t;
      t.c = row[k];
      t.l = has_l ? row[k - 1] & lmask : 0u;
      t.r = has_r ? row[k + 1] & rmask : 0u;
      t.s = t.c & colmask;
      if (MASKED && !row_is_source(rr)) { t.s = 0u; t.l = 0u; t.r = 0u; }
      return t;
    };
#pragma unroll 1
    for (int dir = 0; dir < 2; ++dir) {
      const int step = dir == 0 ? 1 : -1;
      int r = dir == 0 ? ra : rb - 1;
      // software pipeline: rows r-step (behind), r (here), r+step (ahead) live in registers and the
      // row after that is fetched while this one is computed, so a step's critical path is ALU only
      Row3 behind = load_row(r - step), here = load_row(r), ahead = load_row(r + step);
      uint32_t pw = Pp[r * ws + k];
      for (int cnt = ra; cnt < rb; ++cnt, r += step) {
        const Row3 next = load_row(r + 2 * step);   // rows up to +-2 outside the segment exist (halo >= 2)
        const uint32_t pw_next = (cnt + 1 < rb) ? Pp[(r + step) * ws + k] : 0u;
        const uint32_t m = here.c;
        uint32_t vs = behind.s | ahead.s;
        if (CONN8) vs |= (vs << 1) | (vs >> 1) | ((behind.l | ahead.l) >> 31) | ((behind.r | ahead.r) << 31);
        const uint32_t hz = (here.l >> 31) | (here.r << 31);
        const uint32_t cand = m | ((vs | hz) & pw);
        uint32_t nw;
        const bool src_here = row_is_source(r);
        if (MASKED) {
          const uint32_t rowmask = src_here ? colmask : 0u;
          const uint32_t f = hfill(cand & rowmask, pw & rowmask);
          nw = m | f | (pw & ~rowmask & (cand | (f << 1) | (f >> 1)));
        } else {
          nw = hfill(cand, pw);
        }
        if (nw != m) { MROW(r)[k] = nw; item_changed = 1; }
        behind = here;
        behind.c = nw;
        behind.s = (MASKED && !src_here) ? 0u : nw & colmask;
        here = ahead;
        ahead = next;
        pw = pw_next;
      }
    }
    if (item_changed) {
      changed = 1;
      act_next[it] = 1;
      if (has_l) act_next[it - 1] = 1;
      if (has_r) act_next[it + 1] = 1;
      if (sgi > 0) act_next[it - wpr] = 1;
      if (sgi < nseg - 1) act_next[it + wpr] = 1;
      if (CONN8) {
        if (sgi > 0) { if (has_l) act_next[it - wpr - 1] = 1; if (has_r) act_next[it - wpr + 1] = 1; }
        if (sgi < nseg - 1) { if (has_l) act_next[it + wpr - 1] = 1; if (has_r) act_next[it + wpr + 1] = 1; }
      }
    }
  }
#undef MROW
  return changed;
}

__global__ void __launch_bounds__(SEG_THREADS, 1) k23_segment(const __grid_constant__ SegParams p) {
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ __align__(16) uint32_t smem[];
  const int wpr = p.wpr, ws = seg_row_stride(p.wpr), rows = p.rows_per, C = p.cluster;
  const int rank = C > 1 ? (int)cluster.block_rank() : 0;
  const int slice = blockIdx.x / C;
  const int tid = threadIdx.x;
  // planes: predicate rows owned by this CTA; mask rows with SEG_HALO halo rows above and below
  uint32_t* Pp = smem;                                   // [rows][wpr]
  uint32_t* Mm = smem + (size_t)rows * ws;               // [rows + 2*HALO][ws]  (ws = odd row stride)
  const int nseg_max = (rows + p.seg_rows - 1) / p.seg_rows;
  uint8_t* act0 = reinterpret_cast<uint8_t*>(Mm + (size_t)(rows + 2 * SEG_HALO) * ws);   // [nseg][wpr] x 2
  uint8_t* act1 = act0 + (size_t)nseg_max * wpr;
  __shared__ int s_flag[3];                              // rotating "somebody changed" flags (see below)
#define MROW(r) (Mm + (size_t)((r) + SEG_HALO) * ws)
  const int y_begin = rank * rows;                       // first image row owned
  const int nrows = max(0, min(rows, p.h - y_begin));    // rows actually inside the image
  const uint32_t* gp = p.pred + (size_t)slice * p.h * wpr;
  const uint32_t lastmask = (p.w & 31) ? ((1u << (p.w & 31)) - 1u) : 0xffffffffu;

  for (int i = tid; i < (rows + 2 * SEG_HALO) * ws; i += SEG_THREADS) Mm[i] = 0u;
  for (int i = tid; i < 2 * nseg_max * wpr; i += SEG_THREADS) act0[i] = 0;
  // Launched with programmatic stream serialization: everything above overlaps the tail of the kernel
  // that produces the predicate plane; its memory is only touched after this point.
  cudaGridDependencySynchronize();
  for (int i = tid; i < rows * wpr; i += SEG_THREADS) {
    const int r = i / wpr, k = i - r * wpr;
    uint32_t v = r < nrows ? __ldcg(gp + (size_t)(y_begin + r) * wpr + k) : 0u;
    if (k == wpr - 1) v &= lastmask;
    Pp[r * ws + k] = v;
  }
  if (tid < 3) s_flag[tid] = 0;
  __syncthreads();

  int passes = 0;
  if (p.do_fill) {
    // ---- seeds ------------------------------------------------------------------------------
    // a seed is a change nobody has looked at yet: wake its item AND the items around it (the growth
    // may start in a neighbouring segment or word column although the seed's own item stays as it is)
    auto wake = [&](int r, int k) {
      const int nsg = (nrows + p.seg_rows - 1) / p.seg_rows, sg = r / p.seg_rows;
      for (int a = max(0, sg - 1); a <= min(nsg - 1, sg + 1); ++a)
        for (int b = max(0, k - 1); b <= min(wpr - 1, k + 1); ++b) act0[a * wpr + b] = 1;
    };
    for (int i = tid; i < p.n_seeds; i += SEG_THREADS) {
      const int sx = p.seeds[i].x, sy = p.seeds[i].y;
      if (!p.host_mode) {
        const int r = sy - y_begin;
        if (r >= 0 && r < nrows) {
          atomicOr(&MROW(r)[sx >> 5], (1u << (sx & 31)) & Pp[r * ws + (sx >> 5)]);
          wake(r, sx >> 5);
        }
      } else {
        // FAST executeOnHost: the seed's neighbours are tested, not the seed itself
        const int ddx[8] = {0, 1, -1, 0, 1, -1, 1, -1}, ddy[8] = {1, 0, 0, -1, 1, 1, -1, -1};
        for (int d = 0; d < (p.conn8 ? 8 : 4); ++d) {
          const int nx = sx + ddx[d], ny = sy + ddy[d];
          const int r = ny - y_begin;
          if (nx < 0 || nx >= p.w || r < 0 || r >= nrows) continue;
          atomicOr(&MROW(r)[nx >> 5], (1u << (nx & 31)) & Pp[r * ws + (nx >> 5)]);
          wake(r, nx >> 5);
        }
      }
    }
    if (C > 1) cluster.sync();  // every CTA's planes and flags are initialised before anyone touches them
    else __syncthreads();

    // ---- fill: sweep passes until nothing changes anywhere in the cluster --------------------
    // Cluster protocol, one cluster.sync per pass: a CTA that changed something in pass q stores 1
    // into flag[q % 3] of EVERY CTA (distributed shared memory); after the sync all CTAs read the same
    // OR.  flag[(q+2) % 3] is cleared for later use — it is next written in pass q+2, which no CTA can
    // reach before this CTA has passed the sync of pass q+1, i.e. after the clear.
    const bool masked = !p.border_expand && !p.host_mode;
    int hchg = 0;
    for (int q = 0;; ++q) {
      ++passes;
      int changed;
      uint8_t* an = (q & 1) ? act1 : act0;
      uint8_t* ax = (q & 1) ? act0 : act1;
      if (p.conn8) changed = masked ? fill_pass<true, true>(Mm, Pp, p, nrows, y_begin, tid, an, ax) : fill_pass<true, false>(Mm, Pp, p, nrows, y_begin, tid, an, ax);
      else changed = masked ? fill_pass<false, true>(Mm, Pp, p, nrows, y_begin, tid, an, ax) : fill_pass<false, false>(Mm, Pp, p, nrows, y_begin, tid, an, ax);
      // the seeds themselves are progress the neighbouring CTAs have not seen: the first round always
      // continues, so that every halo row is refreshed (and its readers woken) at least once
      changed = __syncthreads_or(changed | hchg) | (q == 0 && C > 1);
      if (C == 1) {
        if (!changed) break;
        continue;
      }
      if (changed && tid < C) *cluster.map_shared_rank(&s_flag[q % 3], tid) = 1;
      cluster.sync();
      const int any = s_flag[q % 3];
      if (tid == 0) s_flag[(q + 2) % 3] = 0;
      if (!any) break;
      // refresh the one halo row each side from the neighbouring CTAs; a change there is progress
      hchg = 0;
      for (int k = tid; k < 2 * wpr; k += SEG_THREADS) {
        const int side = k / wpr, kk = k - side * wpr;
        const int nb = side == 0 ? rank - 1 : rank + 1;
        if (nb < 0 || nb >= C) continue;
        const uint32_t* remote = cluster.map_shared_rank(Mm, nb);
        const uint32_t v = remote[(size_t)((side == 0 ? rows - 1 : 0) + SEG_HALO) * ws + kk];
        uint32_t* mine = MROW(side == 0 ? -1 : rows) + kk;
        if (*mine != v) {
          *mine = v;
          hchg = 1;
          // the boundary segment next to that halo row has new input: wake it (and its word neighbours)
          const int nsg = (nrows + p.seg_rows - 1) / p.seg_rows;
          if (nsg > 0) {
            uint8_t* row = ax + (side == 0 ? 0 : nsg - 1) * wpr;
            row[kk] = 1;
            if (kk > 0) row[kk - 1] = 1;
            if (kk < wpr - 1) row[kk + 1] = 1;
          }
        }
      }
      __syncthreads();
    }
  } else {
    for (int i = tid; i < rows * wpr; i += SEG_THREADS) { const int r = i / wpr, k = i - r * wpr; MROW(r)[k] = Pp[r * ws + k]; }
    __syncthreads();
  }
  if (p.iterations && rank == 0 && tid == 0) p.iterations[slice] = passes;

  // ---- halo rows of the final mask for the structuring element --------------------------------
  const int nmax = max((p.dilated || p.dilated_bits) ? p.dil_n : 0, p.eroded ? p.ero_n : 0);
  if (C > 1) {
    cluster.sync();  // all masks final (also orders the do_fill = 0 copy)
    for (int i = tid; i < 2 * nmax * wpr; i += SEG_THREADS) {
      const int side = i / (nmax * wpr), j = i - side * nmax * wpr;
      const int hr = j / wpr, kk = j - hr * wpr;
      const int nb = side == 0 ? rank - 1 : rank + 1;
      if (nb < 0 || nb >= C) continue;
      const uint32_t* remote = cluster.map_shared_rank(Mm, nb);
      if (side == 0) MROW(-1 - hr)[kk] = remote[(size_t)(rows - 1 - hr + SEG_HALO) * ws + kk];
      else MROW(rows + hr)[kk] = remote[(size_t)(hr + SEG_HALO) * ws + kk];
    }
    cluster.sync();  // nobody exits (or overwrites) while a neighbour still reads its rows
  } else {
    __syncthreads();
  }

  // ---- ImageCaster is the identity; dilation / erosion with the row-run form of the SE ---------
  const bool vec = (p.w & 15) == 0 && ((((size_t)p.dilated) | ((size_t)p.eroded) | ((size_t)p.rg_mask)) & 15) == 0;
  for (int i = tid; i < nrows * wpr; i += SEG_THREADS) {
    const int r = i / wpr, k = i - r * wpr;
    const int y = y_begin + r;
    const size_t rowoff = ((size_t)slice * p.h + y) * p.w;
    const uint32_t m = MROW(r)[k];
    const bool has_l = k > 0, has_r = k < wpr - 1;
    if (p.rg_mask) store_word_u8(p.rg_mask + rowoff, k * 32, p.w, m, vec);
    if (p.dilated || p.dilated_bits) {
      uint32_t acc = m;
      if (p.dil_n == 1 && p.dil_hw[0] == 0 && p.dil_hw[1] == 1) {   // Dilation(3) with the disc SE: the 5-pixel plus
        const uint32_t* row = MROW(r);
        acc = m | __funnelshift_r(m, has_r ? row[k + 1] : 0u, 1) | __funnelshift_l(has_l ? row[k - 1] : 0u, m, 1);
        if (y > 0) acc |= MROW(r - 1)[k];
        if (y < p.h - 1) acc |= MROW(r + 1)[k];
      } else if (p.dil_n > 0) {
        acc = 0u;
        for (int b = -p.dil_n; b <= p.dil_n; ++b) {
          const int yy = y + b;
          if (yy < 0 || yy >= p.h) continue;
          const uint32_t* row = MROW(r + b);
          const uint32_t c = row[k], l = has_l ? row[k - 1] : 0u, rr = has_r ? row[k + 1] : 0u;
          const int hw = p.dil_hw[b + p.dil_n];
          acc |= c;
          for (int s2 = 1; s2 <= hw; ++s2) acc |= __funnelshift_r(c, rr, s2) | __funnelshift_l(l, c, s2);
        }
      }
      if (p.dilated) store_word_u8(p.dilated + rowoff, k * 32, p.w, acc, vec);
      if (p.dilated_bits)   // dilation may have set bits beyond the last pixel of the row: they are not part of the image
        p.dilated_bits[((size_t)slice * p.h + y) * wpr + k] = (k == wpr - 1 && (p.w & 31)) ? acc & ((1u << (p.w & 31)) - 1u) : acc;
    }
    if (p.eroded) {
      uint32_t acc = m;
      if (p.ero_n > 0) {
        for (int b = -p.ero_n; b <= p.ero_n && acc; ++b) {
          const int yy = y + b;
          if (yy < 0 || yy >= p.h) { acc = 0u; break; }
          const uint32_t* row = MROW(r + b);
          const uint32_t c = row[k], l = has_l ? row[k - 1] : 0u, rr = has_r ? row[k + 1] : 0u;
          const int hw = p.ero_hw[b + p.ero_n];
          acc &= c;
          for (int s2 = 1; s2 <= hw; ++s2) acc &= __funnelshift_r(c, rr, s2) & __funnelshift_l(l, c, s2);
        }
      }
      store_word_u8(p.eroded + rowoff, k * 32, p.w, acc, vec);
    }
  }
#undef MROW
}

// ---- global-memory fallback for slices that exceed one cluster's shared memory ------------------------
// The mask lives bit-packed in global memory.  One CTA takes a band of rows, loads it (plus one halo
// row each side) into shared memory, grows it to its local fixed point with the same sweep passes as
// K23, and writes it back; information crosses a band boundary once per launch.  The host relaunches
// until a launch reports no change through the global flag.
__global__ void __launch_bounds__(SEG_THREADS, 1) k2_band_fill(const __grid_constant__ SegParams p, uint32_t* __restrict__ gmask,
                                                               int* __restrict__ gflag, int first_launch) {
  extern __shared__ __align__(16) uint32_t smem[];
  const int wpr = p.wpr, ws = seg_row_stride(p.wpr), rows = p.rows_per;
  const int nbands = (p.h + rows - 1) / rows;
  const int slice = blockIdx.x / nbands, band = blockIdx.x - slice * nbands;
  const int tid = threadIdx.x;
  uint32_t* Pp = smem;
  uint32_t* Mm = smem + (size_t)rows * ws;
  const int nseg_max = (rows + p.seg_rows - 1) / p.seg_rows;
  uint8_t* act0 = reinterpret_cast<uint8_t*>(Mm + (size_t)(rows + 2 * SEG_HALO) * ws);
  uint8_t* act1 = act0 + (size_t)nseg_max * wpr;
#define MROW(r) (Mm + (size_t)((r) + SEG_HALO) * ws)
  const int y_begin = band * rows;
  const int nrows = min(rows, p.h - y_begin);
  const uint32_t* gp = p.pred + (size_t)slice * p.h * wpr;
  uint32_t* gm = gmask + (size_t)slice * p.h * wpr;
  const uint32_t lastmask = (p.w & 31) ? ((1u << (p.w & 31)) - 1u) : 0xffffffffu;
  for (int i = tid; i < (rows + 2 * SEG_HALO) * ws; i += SEG_THREADS) Mm[i] = 0u;
  for (int i = tid; i < 2 * nseg_max * wpr; i += SEG_THREADS) act0[i] = 1;   // everything is looked at once per launch
  __syncthreads();
  for (int i = tid; i < rows * wpr; i += SEG_THREADS) {
    const int r = i / wpr, k = i - r * wpr;
    uint32_t v = r < nrows ? gp[(size_t)(y_begin + r) * wpr + k] : 0u;
    if (k == wpr - 1) v &= lastmask;
    Pp[r * ws + k] = v;
    if (!first_launch && r < nrows) MROW(r)[k] = gm[(size_t)(y_begin + r) * wpr + k];
  }
  if (!first_launch)
    for (int k = tid; k < 2 * wpr; k += SEG_THREADS) {   // halo rows written by the neighbouring bands in earlier launches
      const int side = k / wpr, kk = k - side * wpr;
      const int y = side == 0 ? y_begin - 1 : y_begin + nrows;
      if (y >= 0 && y < p.h) MROW(side == 0 ? -1 : nrows)[kk] = gm[(size_t)y * wpr + kk];
    }
  __syncthreads();
  if (first_launch) {
    for (int i = tid; i < p.n_seeds; i += SEG_THREADS) {
      const int sx = p.seeds[i].x, sy = p.seeds[i].y;
      if (!p.host_mode) {
        const int r = sy - y_begin;
        if (r >= 0 && r < nrows) atomicOr(&MROW(r)[sx >> 5], (1u << (sx & 31)) & Pp[r * ws + (sx >> 5)]);
      } else {
        const int ddx[8] = {0, 1, -1, 0, 1, -1, 1, -1}, ddy[8] = {1, 0, 0, -1, 1, 1, -1, -1};
        for (int d = 0; d < (p.conn8 ? 8 : 4); ++d) {
          const int nx = sx + ddx[d], ny = sy + ddy[d];
          const int r = ny - y_begin;
          if (nx < 0 || nx >= p.w || r < 0 || r >= nrows) continue;
          atomicOr(&MROW(r)[nx >> 5], (1u << (nx & 31)) & Pp[r * ws + (nx >> 5)]);
        }
      }
    }
    __syncthreads();
  }
  const bool masked = !p.border_expand && !p.host_mode;
  int ever = first_launch;   // the seeds themselves are a change
  for (int q = 0;; ++q) {
    uint8_t* an = (q & 1) ? act1 : act0;
    uint8_t* ax = (q & 1) ? act0 : act1;
    int changed;
    if (p.conn8) changed = masked ? fill_pass<true, true>(Mm, Pp, p, nrows, y_begin, tid, an, ax) : fill_pass<true, false>(Mm, Pp, p, nrows, y_begin, tid, an, ax);
    else changed = masked ? fill_pass<false, true>(Mm, Pp, p, nrows, y_begin, tid, an, ax) : fill_pass<false, false>(Mm, Pp, p, nrows, y_begin, tid, an, ax);
    changed = __syncthreads_or(changed);
    if (!changed) break;
    ever = 1;
  }
  if (ever && !first_launch && tid == 0) *gflag = 1;
  if (first_launch && tid == 0) *gflag = 1;           // a second launch always runs so that the halos are seen once
  for (int i = tid; i < nrows * wpr; i += SEG_THREADS) {
    const int r = i / wpr, k = i - r * wpr;
    gm[(size_t)(y_begin + r) * wpr + k] = MROW(r)[k];
  }
#undef MROW
}

// cast + dilation / erosion from the global bit-packed mask (same row-run form as in K23)
__global__ void k3_global_morph(const __grid_constant__ SegParams p, const uint32_t* __restrict__ gmask, int passes) {
  const int wpr = p.wpr;
  const long long total = (long long)p.n * p.h * wpr;
  const bool vec = (p.w & 15) == 0 && ((((size_t)p.dilated) | ((size_t)p.eroded) | ((size_t)p.rg_mask)) & 15) == 0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i % wpr);
    const long long row = i / wpr;
    const int y = (int)(row % p.h), slice = (int)(row / p.h);
    if (p.iterations && k == 0 && y == 0) p.iterations[slice] = passes;
    const uint32_t* gm = gmask + (size_t)slice * p.h * wpr;
    auto word = [&](int yy, int kk) -> uint32_t { return (yy < 0 || yy >= p.h || kk < 0 || kk >= wpr) ? 0u : gm[(size_t)yy * wpr + kk]; };
    const size_t rowoff = ((size_t)slice * p.h + y) * p.w;
    const uint32_t m = word(y, k);
    if (p.rg_mask) store_word_u8(p.rg_mask + rowoff, k * 32, p.w, m, vec);
    if (p.dilated || p.dilated_bits) {
      uint32_t acc = p.dil_n > 0 ? 0u : m;
      for (int b = -p.dil_n; b <= p.dil_n && p.dil_n > 0; ++b) {
        if (y + b < 0 || y + b >= p.h) continue;
        const uint32_t c = word(y + b, k), l = word(y + b, k - 1), rr = word(y + b, k + 1);
        const int hw = p.dil_hw[b + p.dil_n];
        acc |= c;
        for (int s2 = 1; s2 <= hw; ++s2) acc |= __funnelshift_r(c, rr, s2) | __funnelshift_l(l, c, s2);
      }
      if (p.dilated) store_word_u8(p.dilated + rowoff, k * 32, p.w, acc, vec);
      if (p.dilated_bits)
        p.dilated_bits[((size_t)slice * p.h + y) * wpr + k] = (k == wpr - 1 && (p.w & 31)) ? acc & ((1u << (p.w & 31)) - 1u) : acc;
    }
    if (p.eroded) {
      uint32_t acc = m;
      for (int b = -p.ero_n; b <= p.ero_n && p.ero_n > 0 && acc; ++b) {
        if (y + b < 0 || y + b >= p.h) { acc = 0u; break; }
        const uint32_t c = word(y + b, k), l = word(y + b, k - 1), rr = word(y + b, k + 1);
        const int hw = p.ero_hw[b + p.ero_n];
        acc &= c;
        for (int s2 = 1; s2 <= hw; ++s2) acc &= __funnelshift_r(c, rr, s2) & __funnelshift_l(l, c, s2);
      }
      store_word_u8(p.eroded + rowoff, k * 32, p.w, acc, vec);
    }
  }
}

cudaError_t launch_band_fill(const SegParams& p, uint32_t* gmask, int* gflag, int first_launch, cudaStream_t s) {
  const int nbands = (p.h + p.rows_per - 1) / p.rows_per;
  k2_band_fill<<<p.n * nbands, SEG_THREADS, seg_smem_bytes(p.rows_per, p.wpr), s>>>(p, gmask, gflag, first_launch);
  return cudaGetLastError();
}

cudaError_t launch_global_morph(const SegParams& p, const uint32_t* gmask, int passes, cudaStream_t s) {
  const long long total = (long long)p.n * p.h * p.wpr;
  long long g = (total + 255) / 256;
  k3_global_morph<<<(int)(g > 148 * 8 ? 148 * 8 : (g < 1 ? 1 : g)), 256, 0, s>>>(p, gmask, passes);
  return cudaGetLastError();
}

size_t seg_smem_bytes(int rows_per, int wpr) {
  // predicate + mask planes (odd row stride), plus two activity-flag arrays (at most one byte per row and word column)
  return ((size_t)rows_per + (size_t)rows_per + 2 * SEG_HALO) * seg_row_stride(wpr) * 4 + 2 * (size_t)rows_per * wpr;
}

cudaError_t launch_segment(const SegParams& p, cudaStream_t s) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(p.n * p.cluster);
  cfg.blockDim = dim3(SEG_THREADS);
  cfg.dynamicSmemBytes = seg_smem_bytes(p.rows_per, p.wpr);
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = p.cluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  return cudaLaunchKernelEx(&cfg, k23_segment, p);
}

// =================================================================================================
// helpers for the per-operator entry points
// =================================================================================================
__global__ void k_threshold_pack(const float* __restrict__ img, int n, int w, int h, float lo, float hi,
                                 uint32_t* __restrict__ bits) {
  const int wpr = (w + 31) >> 5;
  const long long total = (long long)n * h * wpr * 32;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long word = i >> 5;
    const int x = (int)(word % wpr) * 32 + (int)(i & 31);
    const long long row = word / wpr;
    bool in = false;
    if (x < w) { const float v = img[row * w + x]; in = v >= lo && v <= hi; }
    const uint32_t b = __ballot_sync(0xffffffffu, in);
    if ((i & 31) == 0) bits[word] = b;
  }
}

__global__ void k_pack_u8(const uint8_t* __restrict__ mask, int n, int w, int h, uint32_t* __restrict__ bits) {
  const int wpr = (w + 31) >> 5;
  const long long total = (long long)n * h * wpr * 32;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long word = i >> 5;
    const int x = (int)(word % wpr) * 32 + (int)(i & 31);
    const long long row = word / wpr;
    const bool in = x < w && mask[row * w + x] == 1;
    const uint32_t b = __ballot_sync(0xffffffffu, in);
    if ((i & 31) == 0) bits[word] = b;
  }
}

static int grid_for(long long total) {
  long long g = (total + 255) / 256;
  return (int)(g < 1 ? 1 : (g > 148 * 16 ? 148 * 16 : g));
}

cudaError_t launch_threshold_pack(const float* img, int n, int w, int h, float lo, float hi, uint32_t* bits, cudaStream_t s) {
  const long long total = (long long)n * h * ((w + 31) >> 5) * 32;
  k_threshold_pack<<<grid_for(total), 256, 0, s>>>(img, n, w, h, lo, hi, bits);
  return cudaGetLastError();
}

cudaError_t launch_pack_u8(const uint8_t* mask, int n, int w, int h, uint32_t* bits, cudaStream_t s) {
  const long long total = (long long)n * h * ((w + 31) >> 5) * 32;
  k_pack_u8<<<grid_for(total), 256, 0, s>>>(mask, n, w, h, bits);
  return cudaGetLastError();
}

cudaError_t configure_kernels() {
  cudaError_t e;
#define NM03_CFG(k) if ((e = config_k1_w##k()) != cudaSuccess) return e;
  NM03_K1_WINDOWS(NM03_CFG)
#undef NM03_CFG
  e = cudaFuncSetAttribute(k23_segment, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 64);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k2_band_fill, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 64);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(k23_segment, cudaFuncAttributeNonPortableClusterSizeAllowed, 0);
}

}  // namespace nm03
