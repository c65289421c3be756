// Included by hands.cu only: the kernel body and its host launcher.
namespace ag2 {
namespace {

// Row spans of the query window, tightened to the chord of the sphere in each row.
__device__ void setup_rows(Fixed& sh, const GridDev& g, const float q[3], float rq, int tid) {
  QueryWindow w = make_window(g, q[0], q[1], q[2], rq);
  int rows = min(w.rows(), kMaxRows);
  int ny = w.y1 - w.y0 + 1;
  float h = 1.0f / g.inv_h;
  for (int row = tid; row < kMaxRows; row += kHandThreads) {
    uint32_t cnt = 0, begin = 0;
    if (row < rows) {
      int cz = w.z0 + row / ny, cy = w.y0 + row % ny;
      // distance from q to the cell slab in y and z (clamped edge cells are unbounded)
      float ylo = g.oy + (float)cy * h, yhi = ylo + h;
      float zlo = g.oz + (float)cz * h, zhi = zlo + h;
      float dy = 0.f, dz = 0.f;
      if (q[1] < ylo && cy > 0) dy = ylo - q[1];
      if (q[1] > yhi && cy < g.ny - 1) dy = q[1] - yhi;
      if (q[2] < zlo && cz > 0) dz = zlo - q[2];
      if (q[2] > zhi && cz < g.nz - 1) dz = q[2] - zhi;
      // conservative: shrink the slab distances, grow the radius
      dy = fmaxf(0.f, dy * 0.999f - 1e-5f);
      dz = fmaxf(0.f, dz * 0.999f - 1e-5f);
      float rem = rq * rq - dy * dy - dz * dz;
      if (rem > 0.f) {
        float wx = sqrtf(rem) * 1.001f + 1e-5f;
        int x0 = cell_coord(q[0] - wx, g.ox, g.inv_h, g.nx);
        int x1 = cell_coord(q[0] + wx, g.ox, g.inv_h, g.nx);
        uint32_t c0 = cell_id(g, x0, cy, cz);
        begin = __ldg(&g.cell_start[c0]);
        uint32_t end = __ldg(&g.cell_start[c0 + (uint32_t)(x1 - x0) + 1]);
        cnt = end - begin;
      }
    }
    sh.row_begin[row] = begin;
    sh.row_prefix[row + 1] = cnt;
  }
  if (tid == 0) {
    sh.row_prefix[0] = 0;
    sh.nrows = (uint32_t)rows;
  }
  __syncthreads();
  // inclusive scan of kMaxRows counts by warp 0 (32 lanes x 32 items)
  if (tid < 32) {
    constexpr int per = kMaxRows / 32;
    uint32_t local = 0;
    for (int k = 0; k < per; k++) local += sh.row_prefix[1 + tid * per + k];
    uint32_t incl = local;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(AG2_FULL, incl, o);
      if (tid >= o) incl += t;
    }
    uint32_t run = incl - local;
    for (int k = 0; k < per; k++) {
      run += sh.row_prefix[1 + tid * per + k];
      sh.row_prefix[1 + tid * per + k] = run;
    }
    if (tid == 31) sh.total_cand = run;
  }
  __syncthreads();
  // first cell row of every fill batch: one search per BATCH here instead of one per thread and batch in fill_chunk
  {
    const uint32_t T = sh.total_cand;
    const uint32_t nb = min((uint32_t)kBatchRows, (T + kFillBatch - 1) / kFillBatch);
    for (uint32_t b = tid; b < nb; b += kHandThreads) {
      const uint32_t tb = b * kFillBatch;
      int lo = 0, hi = (int)sh.nrows;  // prefix[lo] <= tb < prefix[hi]
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (sh.row_prefix[mid] <= tb) lo = mid; else hi = mid;
      }
      sh.batch_row[b] = (uint16_t)lo;
    }
  }
  __syncthreads();
}

// Stages the next chunk of cropped neighbours, starting at candidate sh.cursor (a multiple of the batch size, or
// T).  Warps take batches of 32 x kCand consecutive candidates from a shared counter -- no block-wide barrier
// inside the loop.  A warp takes a batch only while (points staged so far) + (everything all warps may still be
// adding) fits the staging buffer: the LAST batch ever taken saw nkept + kInFlight <= cap, and what was not yet
// counted in that nkept is at most one batch per warp, so the buffer cannot overflow; a refused warp stops, and
// since nkept only grows every later request is refused too: the batches taken are a prefix of the sequence.
// Each thread takes kCand CONSECUTIVE candidates: its batch's first row from the table of setup_rows, a short walk to
// its own row, then sequential row advance.
__device__ void fill_chunk(Fixed& sh, float4* cache, const HandsArgs& A, const HandConst& hc,
                           const float q[3], bool count_stats, int tid) {
  const int lane = tid & 31;
  constexpr uint32_t kBatch = kFillBatch;
  constexpr uint32_t kInFlight = (kHandThreads / 32) * kBatch;
  const uint32_t T = sh.total_cand;
  if (tid == 0) {
    sh.nkept = 0;
    sh.next_batch = (sh.cursor + kBatch - 1) / kBatch;
  }
  __syncthreads();
  const double az0 = sh.F[0][2], az1 = sh.F[1][2], az2 = sh.F[2][2];
  const float fz0 = sh.Fz[0], fz1 = sh.Fz[1], fz2 = sh.Fz[2], hh_in = sh.hh_in, hh_out = sh.hh_out;
  uint32_t kcount = 0;
  for (;;) {
    uint32_t batch = 0xFFFFFFFFu;
    if (lane == 0) {
      const uint32_t seen = *reinterpret_cast<volatile uint32_t*>(&sh.nkept);
      if (seen + kInFlight <= A.cap) batch = atomicAdd(&sh.next_batch, 1u);
    }
    batch = __shfl_sync(AG2_FULL, batch, 0);
    if (batch == 0xFFFFFFFFu || (unsigned long long)batch * kBatch >= T) break;
    const uint32_t tb = batch * kBatch + (uint32_t)lane * kCand;
    int row = 0;
    uint32_t row_end = 0, pos = 0;
    if (tb < T) {
      int lo;
      if (batch < (uint32_t)kBatchRows) {  // the batch's first row is tabulated: a lane walks the few rows to its own
        lo = sh.batch_row[batch];
        while (sh.row_prefix[lo + 1] <= tb) lo++;
      } else {
        int hi = (int)sh.nrows;  // prefix[lo] <= tb < prefix[hi]
        lo = 0;
        while (hi - lo > 1) {
          int mid = (lo + hi) >> 1;
          if (sh.row_prefix[mid] <= tb) lo = mid; else hi = mid;
        }
      }
      row = lo;
      row_end = sh.row_prefix[row + 1];
      pos = sh.row_begin[row] + (tb - sh.row_prefix[row]);
    }
    // positions first, then all loads in flight, then the tests
    uint32_t pp[kCand];
    bool val[kCand];
#pragma unroll
    for (int b = 0; b < kCand; b++) {
      const uint32_t t = tb + (uint32_t)b;
      val[b] = t < T;
      pp[b] = 0;
      if (val[b]) {
        if (t >= row_end) {
          do {
            row++;
            row_end = sh.row_prefix[row + 1];
          } while (t >= row_end);
          pos = sh.row_begin[row];
        }
        pp[b] = pos++;
      }
    }
    float4 pv[kCand];
#pragma unroll
    for (int b = 0; b < kCand; b++) pv[b] = val[b] ? A.g.pts[pp[b]] : make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t keepmask = 0;
#pragma unroll
    for (int b = 0; b < kCand; b++) {
      if (val[b]) {
        // radiusSearch used q - p; hand_search.cpp:209 centres with p - q (exact negation)
        const float dx = __fsub_rn(pv[b].x, q[0]);
        const float dy = __fsub_rn(pv[b].y, q[1]);
        const float dz = __fsub_rn(pv[b].z, q[2]);
        const float d2 = dist2_f(dx, dy, dz);
        if (d2 < hc.r2_hands) {
          kcount++;
          // crop -hh < z < hh (hand_search.cpp:334): FP32 first, FP64 only within eps32 of the bounds
          const float z32 = fabsf(fmaf(fz2, dz, fmaf(fz1, dy, fz0 * dx)));
          bool keep = z32 < hh_in;
          if (!keep && !(z32 > hh_out)) {
            const double zf = (az0 * (double)dx + az1 * (double)dy) + az2 * (double)dz;
            keep = zf > hc.neg_hand_height && zf < hc.hand_height;
          }
          if (keep) keepmask |= 1u << b;
        }
        pv[b] = make_float4(dx, dy, dz, __uint_as_float(pp[b]));
      }
    }
    // one warp scan + one shared atomic per warp and batch (the order inside the staging buffer is free:
    // every consumer is an associative reduction or a set)
    const uint32_t mine = (uint32_t)__popc(keepmask);
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t up = __shfl_up_sync(AG2_FULL, incl, o);
      if (lane >= o) incl += up;
    }
    const uint32_t wtot = __shfl_sync(AG2_FULL, incl, 31);
    uint32_t base = 0;
    if (lane == 31 && wtot) base = atomicAdd(&sh.nkept, wtot);
    base = __shfl_sync(AG2_FULL, base, 31);
    uint32_t w = base + incl - mine;
#pragma unroll
    for (int b = 0; b < kCand; b++)
      if (keepmask & (1u << b)) cache[w++] = pv[b];
  }
  __syncthreads();  // every batch taken is staged, nkept is final
  if (tid == 0) sh.cursor = (uint32_t)min((unsigned long long)T, (unsigned long long)sh.next_batch * kBatch);
  if (count_stats) {
    for (int o = 16; o > 0; o >>= 1) kcount += __shfl_xor_sync(AG2_FULL, kcount, o);
    if (lane == 0 && kcount) atomicAdd(&sh.sumK, (unsigned long long)kcount);
  }
  __syncthreads();
  if (count_stats && tid == 0) sh.sumKc += sh.nkept;
  __syncthreads();
}

// 20-slot collision mask of a point at hand-frame x: bit i iff fs_lo[i] < x < fs_hi[i].
__device__ __forceinline__ uint32_t slot_mask(const Fixed& sh, double x, double x0, double xscale) {
  double bf = (x - x0) * xscale;
  int bi = bf > 0.0 ? (bf < (double)(AG2_SLOT_BUCKETS - 1) ? (int)bf : AG2_SLOT_BUCKETS - 1) : 0;
  int k = sh.kstart[bi];
  while (k < AG2_SLOT_EDGES && sh.U[k] < x) k++;
  const bool eq = k < AG2_SLOT_EDGES && sh.U[k] == x;
  return eq ? sh.Meq[k] : sh.M[k];
}

// FP32 filter look-ups.  A bucket entry without the top bit is the answer.  With it, one threshold k is near the
// bucket: certain below lo[k] / above hi[k] (two shared loads, no call); only a coordinate within eps32 of the
// threshold, or a bucket with several thresholds nearby, takes the exact FP64 evaluation.
__device__ __noinline__ int exact_class(const Fixed& sh, const Orient& o, const float4 p, int ND) {
  return depth_class(dot3d_ref(o.FRy, p), sh.depth, ND, 200.0);
}
__device__ __noinline__ uint32_t exact_mask(const Fixed& sh, const Orient& o, const float4 p, double x0, double xscale) {
  return slot_mask(sh, dot3d_ref(o.FRx, p), x0, xscale);
}
__device__ __forceinline__ int class_of(const Fixed& sh, const Orient& o, const float4& p, float yf, uint32_t v, int ND) {
  if (!(v & 0x80000000u)) return (int)v;
  if ((v & 0xFFFFu) != 0xFFFFu) {
    const uint32_t k0 = v & 0xFFu, k1 = (v >> 8) & 0xFFu;
    if (yf < __uint_as_float(sh.lut[kLutYLo + k0])) return (int)k0;
    if (yf > __uint_as_float(sh.lut[kLutYHi + k1])) return (int)k1 + 1;
  }
  return exact_class(sh, o, p, ND);
}
__device__ __forceinline__ uint32_t mask_of(const Fixed& sh, const Orient& o, const float4& p, float xf, uint32_t v,
                                            double x0, double xscale) {
  if (!(v & 0x80000000u)) return v;
  if ((v & 0xFFFFu) != 0xFFFFu) {
    const uint32_t k0 = v & 0xFFu, k1 = (v >> 8) & 0xFFu;
    if (xf < __uint_as_float(sh.lut[kLutXLo + k0])) return sh.M[k0];
    if (xf > __uint_as_float(sh.lut[kLutXHi + k1])) return sh.M[k1 + 1];
  }
  return exact_mask(sh, o, p, x0, xscale);
}
// closing-region membership y < top && left < x < right (finger_hand.cpp:170), FP32 first
__device__ __forceinline__ bool in_box(const Orient& o, const float4& p, float xf, float yf) {
  if (yf < o.top_in && xf > o.left_in && xf < o.right_in) return true;
  if (yf > o.top_out || xf < o.left_out || xf > o.right_out) return false;
  const double x = dot3d_ref(o.FRx, p), y = dot3d_ref(o.FRy, p);
  return y < o.top && x > o.left && x < o.right;
}

__global__ void __launch_bounds__(kHandThreads, 2)
k_hands(HandsArgs A, const __grid_constant__ HandConst hc) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int R = hc.num_orientations, ND = hc.num_depths;
  Fixed& sh = *reinterpret_cast<Fixed*>(smem_raw);
  Orient* ors = reinterpret_cast<Orient*>(smem_raw + ((sizeof(Fixed) + 15) & ~(size_t)15));
  uint32_t* occ = reinterpret_cast<uint32_t*>(
      smem_raw + ((((sizeof(Fixed) + 15) & ~(size_t)15) + (size_t)R * sizeof(Orient) + 15) & ~(size_t)15));
  float4* cache = reinterpret_cast<float4*>(smem_raw + hands_smem_base(R, ND));
  const int tid = threadIdx.x, lane = tid & 31;
  const uint32_t s = blockIdx.x;
  const ag2_frame& fr = A.frames[s];
  // default records
  for (int r = tid; r < R; r += kHandThreads) {
    ag2_sweep_record sw;
    sw.fingers = 0, sw.hand = 0, sw.slot = -1, sw.depth_step = -1, sw.num_in_box = -1;
    sw.antipodal = 0, sw.emitted = 0, sw.reserved = 0;
    A.sweeps[(size_t)s * R + r] = sw;
    A.box_ref[(size_t)s * R + r] = BoxRef{0u, 0u};
  }
  if (!fr.valid) return;
  float q[3] = {(float)fr.sample[0], (float)fr.sample[1], (float)fr.sample[2]};
  if (tid < 9) {
    int a = tid / 3, b = tid % 3;
    sh.F[a][b] = b == 0 ? fr.normal[a] : (b == 1 ? fr.binormal[a] : fr.curvature_axis[a]);
  }
  if (tid < 3) sh.sample[tid] = fr.sample[tid];
  if (tid == 0) {
    sh.cursor = 0;
    sh.sumK = 0;
    sh.sumKc = 0;
    sh.pool_ok = 1;
  }
  // tables: constant bank -> shared (they are indexed per lane)
  for (int i = tid; i < kLutWords; i += kHandThreads) sh.lut[i] = __ldg(&A.lut[i]);
  for (int i = tid; i < AG2_SLOT_EDGES; i += kHandThreads) {
    sh.U[i] = hc.lut.U[i];
    sh.Meq[i] = hc.lut.Meq[i];
  }
  for (int i = tid; i <= AG2_SLOT_EDGES; i += kHandThreads) sh.M[i] = hc.lut.M[i];
  for (int i = tid; i < AG2_SLOT_BUCKETS; i += kHandThreads) sh.kstart[i] = hc.lut.kstart[i];
  for (int i = tid; i < ND; i += kHandThreads) {
    sh.depth[i] = hc.depth[i];
    sh.bottom[i] = hc.bottom[i];
  }
  for (int r = tid; r < R; r += kHandThreads) {
    Orient& o = ors[r];
    o.ymin = d_ordered(INFINITY);
    o.m = 0;
    o.minx = d_ordered(INFINITY);
    o.maxx = d_ordered(-INFINITY);
    o.fill = 0;
    o.anyl = o.anyr = 0;
    for (int k = 0; k < 8; k++) o.ap[k] = d_ordered((k & 1) ? INFINITY : -INFINITY);
    o.feasible = 0;
  }
  for (int i = tid; i < R * ND; i += kHandThreads) occ[i] = 0;
  __syncthreads();
  // frame_rot = frame * rot (hand_search.cpp:357-358), columns 0 and 1; column 2 == frame column 2
  for (int t = tid; t < R * 3; t += kHandThreads) {
    int r = t / 3, a = t % 3;
    double ca = hc.cos_a[r], sa = hc.sin_a[r];
    double nsa = -1.0 * sa;
    ors[r].FRx[a] = (sh.F[a][0] * ca + sh.F[a][1] * sa) + sh.F[a][2] * 0.0;
    ors[r].FRy[a] = (sh.F[a][0] * nsa + sh.F[a][1] * ca) + sh.F[a][2] * 0.0;
    ors[r].fx[a] = (float)ors[r].FRx[a];
    ors[r].fy[a] = (float)ors[r].FRy[a];
  }
  if (tid < 3) sh.Fz[tid] = (float)sh.F[tid][2];
  if (tid == 3) {
    sh.hh_in = nextafterf((float)(hc.hand_height - (double)hc.eps32), -INFINITY);
    sh.hh_out = nextafterf((float)(hc.hand_height + (double)hc.eps32), INFINITY);
  }
  setup_rows(sh, A.g, q, hc.rq_hands, tid);
  const uint32_t T = sh.total_cand;
  const double inv_dstep = 200.0;  // 1/0.005: estimator only, exact fix-up follows
  const double lut_x0 = hc.lut.x0, lut_xs = hc.lut.xscale;
  const float e2 = 2.0f * hc.eps32;
  const float yl0 = hc.ylut_v0, yls = hc.ylut_scale, xl0 = hc.xlut_v0, xls = hc.xlut_scale;

  // ---------------- pass 1: slot occupancy per depth class and min y, per orientation
  bool single_chunk = true;
  for (int chunk = 0;; chunk++) {
    fill_chunk(sh, cache, A, hc, q, true, tid);
    const uint32_t nk = sh.nkept;
    if (chunk == 0 && sh.cursor < T) single_chunk = false;
    for (int r = 0; r < R; r++) {
      Orient& o = ors[r];
      const float fx0 = o.fx[0], fx1 = o.fx[1], fx2 = o.fx[2];
      const float fy0 = o.fy[0], fy1 = o.fy[1], fy2 = o.fy[2];
      uint32_t* occ_r = occ + r * ND;
      // FP32 minimum of y over this thread's points, its position, and whether another point came within 2 eps32
      float m = INFINITY;
      uint32_t imin = 0xFFFFFFFFu;
      bool amb = false;
      // four points per iteration: their loads, dot products and table look-ups are independent of each other
      constexpr int U = 4;
      for (uint32_t i0 = tid; i0 < nk; i0 += U * kHandThreads) {
        float4 p[U];
        float yf[U], xf[U];
        uint32_t vy[U], vx[U];
#pragma unroll
        for (int u = 0; u < U; u++) p[u] = cache[min(i0 + u * kHandThreads, nk - 1)];
#pragma unroll
        for (int u = 0; u < U; u++) {
          yf[u] = dot3f(fy0, fy1, fy2, p[u]);
          xf[u] = dot3f(fx0, fx1, fx2, p[u]);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
          vy[u] = sh.lut[kLutY + lut_bucket(yf[u], yl0, yls, kYLutBuckets)];
          vx[u] = sh.lut[kLutX + lut_bucket(xf[u], xl0, xls, kXLutBuckets)];
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
          const uint32_t i = i0 + u * kHandThreads;
          if (i >= nk) break;
          if (yf[u] < m + e2) {
            amb = !(yf[u] < m - e2);
            if (yf[u] < m) m = yf[u], imin = i;
          }
          const int ct = class_of(sh, o, p[u], yf[u], vy[u], ND);
          if (ct < ND) {
            const uint32_t mask = mask_of(sh, o, p[u], xf[u], vx[u], lut_x0, lut_xs);
            // OR is idempotent and monotone: only touch the word when it would change
            if (mask & ~occ_r[ct]) atomicOr(&occ_r[ct], mask);
          }
        }
      }
      // exact min y (points.row(1).minCoeff(), finger_hand.cpp:158): the point that attains it has a float
      // coordinate within 2 eps32 of the float minimum -- of the whole chunk, hence also of its own warp's points --
      // so only the lanes that close to their warp's float minimum evaluate their candidates in FP64
      float wm = m;
      for (int of = 16; of > 0; of >>= 1) wm = fminf(wm, __shfl_xor_sync(AG2_FULL, wm, of));
      if (imin != 0xFFFFFFFFu && m <= wm + e2) {
        double ymin;
        if (!amb) {
          ymin = dot3d_ref(o.FRy, cache[imin]);
        } else {  // several of this thread's points are that close: all of them, exactly
          ymin = INFINITY;
          for (uint32_t i = tid; i < nk; i += kHandThreads) {
            const float4 q4 = cache[i];
            if (dot3f(fy0, fy1, fy2, q4) <= m + e2) ymin = fmin(ymin, dot3d_ref(o.FRy, q4));
          }
        }
        atomicMin(&o.ymin, d_ordered(ymin));
      }
    }
    __syncthreads();
    if (sh.cursor >= T) break;
  }

  // ---------------- finger logic: one thread per orientation
  if (tid < R) {
    const int r = tid;
    Orient& o = ors[r];
    const uint32_t* occ_r = occ + r * ND;
    // depth classes are monotone in y, so the minima over the points follow from min y:
    // mct = class of min y against the bite depths; mcb = class of min y against the
    // hand-base depths, taken over the points with y below the deepest bite (if any)
    const double ymin = ordered_d(o.ymin);
    const int mct = depth_class(ymin, sh.depth, ND, inv_dstep);
    const int mcb = mct < ND ? depth_class(ymin, sh.bottom, ND, inv_dstep) : ND;
    uint32_t cum = occ_r[0];
    // evaluateFingers(points, init_bite): all false if a point is below the hand base
    // or no point is below the bite (finger_hand.cpp:27-42)
    uint32_t fingers = (mcb <= 0 || mct > 0) ? 0u : (~cum & 0xFFFFFu);
    ag2_sweep_record sw;
    sw.fingers = fingers, sw.hand = 0, sw.slot = -1, sw.depth_step = -1, sw.num_in_box = -1;
    sw.antipodal = 0, sw.emitted = 0, sw.reserved = 0;
    if (__popc(fingers) > 2) {  // hand_search.cpp:366
      uint32_t hand = fingers & (fingers >> 10) & 0x3FFu;
      sw.hand = (uint16_t)hand;
      if (hand) {
        int n = __popc(hand);
        int k = (n + 1) / 2 - 1;  // ceil(n / 2.0) - 1, finger_hand.cpp:112
        uint32_t hm = hand;
        for (int i = 0; i < k; i++) hm &= hm - 1;
        int e = __ffs(hm) - 1;
        uint32_t pair = (1u << e) | (1u << (10 + e));
        int last = 0;
        for (int j = 1; j < ND; j++) {  // deepenHand loop, finger_hand.cpp:122-130
          cum |= occ_r[j];
          uint32_t fj = (mcb <= j || mct > j) ? 0u : (~cum & pair);
          if (__popc(fj) < 2) break;
          last = j;
        }
        o.feasible = 1;
        o.slot = e;
        o.last = last;
        o.top = hc.depth[last];
        o.bottom = hc.bottom[last];
        o.left = hc.fs_hi[e];          // finger_spacing_(idx) + finger_width_
        o.right = hc.fs_lo[10 + e];
        o.center = 0.5 * (o.left + o.right);
        double baseline_const = 0.1;
        o.lc = o.left - 0.5 * (baseline_const - (o.right - o.left));
        o.sx = 1.0 / baseline_const;
        o.sy = 1.0 / (o.top - o.bottom);
        o.sz = 1.0 / (2.0 * hc.hand_height);
        // the closing region in float, shrunk (certainly inside) and grown (certainly outside) by eps32
        const double ep = (double)hc.eps32;
        o.top_in = nextafterf((float)(o.top - ep), -INFINITY);
        o.top_out = nextafterf((float)(o.top + ep), INFINITY);
        o.left_in = nextafterf((float)(o.left + ep), INFINITY);
        o.left_out = nextafterf((float)(o.left - ep), -INFINITY);
        o.right_in = nextafterf((float)(o.right - ep), -INFINITY);
        o.right_out = nextafterf((float)(o.right + ep), INFINITY);
        sw.slot = (int8_t)e;
        sw.depth_step = (int8_t)last;
      }
    }
    A.sweeps[(size_t)s * R + r] = sw;
  }
  __syncthreads();
  int nfeas = 0;
  for (int r = 0; r < R; r++) nfeas += ors[r].feasible;
  if (tid == 0) {
    atomicAdd(&A.stat[0], sh.sumK);
    atomicAdd(&A.stat[1], sh.sumKc);
  }
  if (nfeas == 0) return;

  // ---------------- pass 2: closing-region count and x extent
  if (!single_chunk) {
    __syncthreads();
    if (tid == 0) sh.cursor = 0;
    __syncthreads();
  }
  for (;;) {
    if (!single_chunk) fill_chunk(sh, cache, A, hc, q, false, tid);
    const uint32_t nk = sh.nkept;
    for (int r = 0; r < R; r++) {
      const Orient& o = ors[r];
      if (!o.feasible) continue;
      const float fx0 = o.fx[0], fx1 = o.fx[1], fx2 = o.fx[2];
      const float fy0 = o.fy[0], fy1 = o.fy[1], fy2 = o.fy[2];
      uint32_t cnt = 0;
      double mn = INFINITY, mx = -INFINITY;
      for (uint32_t i = tid; i < nk; i += kHandThreads) {
        const float4 p = cache[i];
        if (in_box(o, p, dot3f(fx0, fx1, fx2, p), dot3f(fy0, fy1, fy2, p))) {
          const double x = dot3d_ref(o.FRx, p);  // the extent ends up in the record: exact
          cnt++;
          mn = fmin(mn, x);
          mx = fmax(mx, x);
        }
      }
      for (int of = 16; of > 0; of >>= 1) cnt += __shfl_xor_sync(AG2_FULL, cnt, of);
      mn = warp_min_d(mn);
      mx = warp_max_d(mx);
      if (lane == 0 && cnt) {
        atomicAdd(&ors[r].m, cnt);
        atomicMin(&ors[r].minx, d_ordered(mn));
        atomicMax(&ors[r].maxx, d_ordered(mx));
      }
    }
    __syncthreads();
    if (single_chunk || sh.cursor >= T) break;
  }
  if (tid == 0) {
    uint32_t total = 0, mmax = 0;
    for (int r = 0; r < R; r++) {
      if (!ors[r].feasible) continue;
      ors[r].off = total;
      total += ors[r].m;
      mmax = max(mmax, ors[r].m);
    }
    // 64-bit cursor: the ids of a fused cloud's closing regions can exceed 2^32 in total; what does not fit the
    // pool (BoxRef offsets are 32 bits) is counted, not written, and the host grows the pool or refuses
    const unsigned long long base =
        total ? atomicAdd(reinterpret_cast<unsigned long long*>(A.pool_counter), (unsigned long long)total) : 0ull;
    if (base + total > A.pool_cap) {
      atomicExch(&A.pool_counter[2], 1u);
      sh.pool_ok = 0;
    }
    for (int r = 0; r < R; r++) ors[r].off += (uint32_t)base;
    atomicMax(&A.pool_counter[3], mmax);
    atomicAdd(&A.stat[2], (unsigned long long)total);
  }
  __syncthreads();

  // ---------------- pass 3: in-box ids to the pool, antipodal statistics
  if (!single_chunk) {
    if (tid == 0) sh.cursor = 0;
    __syncthreads();
  }
  const bool pool_ok = sh.pool_ok != 0;
  for (;;) {
    if (!single_chunk) fill_chunk(sh, cache, A, hc, q, false, tid);
    const uint32_t nk = sh.nkept;
    for (int r = 0; r < R; r++) {
      Orient& o = ors[r];
      if (!o.feasible || o.m == 0) continue;
      const double ax0 = o.FRx[0], ax1 = o.FRx[1], ax2 = o.FRx[2];
      const double ay0 = o.FRy[0], ay1 = o.FRy[1], ay2 = o.FRy[2];
      const double az0 = sh.F[0][2], az1 = sh.F[1][2], az2 = sh.F[2][2];
      const double bottom = o.bottom;
      const double sx = o.sx, sy = o.sy, sz = o.sz, lc = o.lc;
      const uint32_t off = o.off;
      // extremal thresholds on SCALED x (antipodal.cpp:16-17; scaling is monotone)
      const double lthr = sx * (ordered_d(o.minx) - lc) + 0.003;
      const double rthr = sx * (ordered_d(o.maxx) - lc) - 0.003;
      bool al = false, ar = false;
      double st[8] = {-INFINITY, INFINITY, -INFINITY, INFINITY, -INFINITY, INFINITY, -INFINITY, INFINITY};
      for (uint32_t base_i = 0; base_i < nk; base_i += kHandThreads) {
        uint32_t i = base_i + tid;
        bool inbox = false;
        double x = 0, y = 0, px = 0, py = 0, pz = 0;
        uint32_t pos = 0;
        if (i < nk) {
          const float4 p = cache[i];
          pos = __float_as_uint(p.w);
          inbox = in_box(o, p, dot3f(o.fx[0], o.fx[1], o.fx[2], p), dot3f(o.fy[0], o.fy[1], o.fy[2], p));
          if (inbox) {
            px = (double)p.x, py = (double)p.y, pz = (double)p.z;
            x = (ax0 * px + ax1 * py) + ax2 * pz;
            y = (ay0 * px + ay1 * py) + ay2 * pz;
          }
        }
        uint32_t ballot = __ballot_sync(AG2_FULL, inbox);
        uint32_t wbase = 0;
        if (lane == 0 && ballot) wbase = atomicAdd(&o.fill, (uint32_t)__popc(ballot));
        wbase = __shfl_sync(AG2_FULL, wbase, 0);
        if (inbox) {
          if (pool_ok) A.pool[off + wbase + __popc(ballot & ((1u << lane) - 1))] = (int32_t)pos;
          double z = (az0 * px + az1 * py) + az2 * pz;
          double xs = sx * (x - lc);
          double ys = sy * (y - bottom);
          double zs = sz * (z - hc.neg_hand_height);
          float4 nv = A.g.nrm[pos];
          double n0 = (double)nv.x, n1 = (double)nv.y, n2 = (double)nv.z;
          double nxr = (ax0 * n0 + ax1 * n1) + ax2 * n2;
          double nyr = (ay0 * n0 + ay1 * n1) + ay2 * n2;
          double nzr = (az0 * n0 + az1 * n1) + az2 * n2;
          double ln = (-1.0 * nxr + 0.0 * nyr) + 0.0 * nzr;  // l^T n, antipodal.cpp:21,24
          double rn = (1.0 * nxr + 0.0 * nyr) + 0.0 * nzr;
          if (ln > hc.cos_friction && xs < lthr) {
            al = true;
            st[0] = fmax(st[0], ys), st[1] = fmin(st[1], ys);
            st[2] = fmax(st[2], zs), st[3] = fmin(st[3], zs);
          }
          if (rn > hc.cos_friction && xs > rthr) {
            ar = true;
            st[4] = fmax(st[4], ys), st[5] = fmin(st[5], ys);
            st[6] = fmax(st[6], zs), st[7] = fmin(st[7], zs);
          }
        }
      }
      uint32_t bl = __ballot_sync(AG2_FULL, al), br = __ballot_sync(AG2_FULL, ar);
      if (bl | br) {  // warp-uniform
#pragma unroll
        for (int k = 0; k < 8; k++) st[k] = (k & 1) ? warp_min_d(st[k]) : warp_max_d(st[k]);
        if (lane == 0) {
          if (bl) {
            o.anyl = 1;
            for (int k = 0; k < 4; k++) {
              if (k & 1) atomicMin(&o.ap[k], d_ordered(st[k])); else atomicMax(&o.ap[k], d_ordered(st[k]));
            }
          }
          if (br) {
            o.anyr = 1;
            for (int k = 4; k < 8; k++) {
              if (k & 1) atomicMin(&o.ap[k], d_ordered(st[k])); else atomicMax(&o.ap[k], d_ordered(st[k]));
            }
          }
        }
      }
    }
    __syncthreads();
    if (single_chunk || sh.cursor >= T) break;
  }

  // ---------------- records: one thread per orientation
  if (tid < R && ors[tid].feasible) {
    const int r = tid;
    const Orient& o = ors[r];
    const size_t slot_id = (size_t)s * R + r;
    ag2_sweep_record sw = A.sweeps[slot_id];
    sw.num_in_box = (int32_t)o.m;
    if (o.m > 0) {  // empty closing region: orientation skipped (hand_search.cpp:377-381)
      int result = 0;
      if (o.anyl || o.anyr) result = 1;
      if (o.anyl && o.anyr) {
        double top_y = fmin(ordered_d(o.ap[0]), ordered_d(o.ap[4]));
        double bot_y = fmax(ordered_d(o.ap[1]), ordered_d(o.ap[5]));
        double top_z = fmin(ordered_d(o.ap[2]), ordered_d(o.ap[6]));
        double bot_z = fmax(ordered_d(o.ap[3]), ordered_d(o.ap[7]));
        if (top_y > bot_y && top_z > bot_z) result = 2;
      }
      ag2_hypothesis g;
      const double surface = ordered_d(o.ymin);  // points.row(1).minCoeff(), finger_hand.cpp:158
      for (int a = 0; a < 3; a++) {
        double fx = o.FRx[a], fy = o.FRy[a], fz = sh.F[a][2];
        g.surface[a] = ((fx * o.center + fy * surface) + fz * 0.0) + sh.sample[a];
        g.bottom[a] = ((fx * o.center + fy * o.bottom) + fz * 0.0) + sh.sample[a];
        g.top[a] = ((fx * o.center + fy * o.top) + fz * 0.0) + sh.sample[a];
        g.binormal[a] = fx;
        g.approach[a] = fy;
        g.axis[a] = fz;
      }
      g.width = ordered_d(o.maxx) - ordered_d(o.minx);
      g.score = 0.0;
      g.sample_id = (int32_t)s;
      g.orientation_id = r;
      g.half_antipodal = (uint8_t)(result >= 1);
      g.full_antipodal = (uint8_t)(result == 2);
      g.finger_slot = (uint8_t)o.slot;
      g.depth_step = (uint8_t)o.last;
      g.num_in_box = (int32_t)o.m;
      g.hand_top = o.top;
      g.hand_bottom = o.bottom;
      A.slots[slot_id] = g;
      A.box_ref[slot_id] = BoxRef{o.off, o.m};
      sw.antipodal = (uint8_t)result;
      sw.emitted = 1;
    }
    A.sweeps[slot_id] = sw;
  }
}

}  // namespace

int hands_launch(ag2_ctx* c) {
  phase_begin(c, 3);
  cudaStream_t st = c->stream;
  const uint32_t S = (uint32_t)c->num_samples;
  const int R = c->p.num_orientations;
  const size_t SR = (size_t)S * R;
  AG2_CUDA(c, c->sweeps.reserve(SR));
  AG2_CUDA(c, c->slots.reserve(SR));
  AG2_CUDA(c, c->box_ref.reserve(SR));
  if (c->box_pool.cap == 0) AG2_CUDA(c, c->box_pool.reserve(std::max<size_t>((size_t)1 << 22, SR * 256)));
  // two CTAs per SM: staging capacity = what is left of half the shared memory
  const size_t base = hands_smem_base(R, c->hc.num_depths);
  size_t budget = 112 * 1024;
  if (const char* e = getenv("AG2_HANDS_BUDGET_KB")) budget = (size_t)std::min(224L, std::max(64L, atol(e))) * 1024;  // measurement hook
  if (base + (size_t)(kHandThreads * kCand + 256) * sizeof(float4) > budget) {
    c->err = "hands: num_orientations x deepen steps too large for shared memory";
    return AG2_E_CAPACITY;
  }
  uint32_t cap = (uint32_t)(((budget - base) / sizeof(float4)) & ~(size_t)127);
  // test hook: a small staging capacity forces the chunked (multi-pass) path on small scenes
  if (const char* e = getenv("AG2_HANDS_STAGE_CAP")) {
    long v = atol(e);
    if (v >= kHandThreads * kCand + 128 && (uint32_t)v < cap) cap = (uint32_t)v & ~127u;
  }
  const size_t smem = base + (size_t)cap * sizeof(float4);
  AG2_CUDA(c, cudaFuncSetAttribute(k_hands, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  AG2_CUDA(c, cudaMemsetAsync(c->scalars.p + 20, 0, 12 * sizeof(uint32_t), st));
  HandsArgs A;
  A.g = grid_dev(c);
  A.lut = c->hands_lut.p;
  A.frames = c->frames.p;
  A.cap = cap;
  A.sweeps = c->sweeps.p;
  A.slots = c->slots.p;
  A.box_ref = c->box_ref.p;
  A.pool = c->box_pool.p;
  A.pool_cap = (uint32_t)std::min<size_t>(c->box_pool.cap, 0xFFFFFFFFu);
  A.pool_counter = c->scalars.p + 20;  // [0..1] 64-bit cursor, [2] overflow flag, [3] max m
  A.stat = (unsigned long long*)(c->scalars.p + 24);
  k_hands<<<S, kHandThreads, smem, st>>>(A, c->hc);
  AG2_LAUNCH_CHECK(c);
  // cursor, overflow flag, statistics: to pinned memory behind the kernel; hands_finish reads them after the
  // next synchronisation (ag2_detect has ONE for the sweep, the two lists and the prune step together)
  AG2_CUDA(c, cudaMemcpyAsync(c->h_pin + 8, c->scalars.p + 20, 12 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaMemcpyAsync(c->h_pin + 20, c->scalars.p + 13, 4, cudaMemcpyDeviceToHost, st));
  phase_end(c, 3);
  return AG2_OK;
}

// After a synchronisation: AG2_OK, kHandsRetry (the in-box pool was too small and has been grown: launch again),
// or an error.
int hands_finish(ag2_ctx* c) {
  const uint32_t* res = c->h_pin + 8;
  unsigned long long need, stat[3];
  memcpy(&need, res, 8);
  memcpy(stat, res + 4, sizeof(stat));
  if (res[2]) {  // pool overflow: grow to the exact need and redo
    if (need > 0xFFFFFFF0ull) {
      c->err = "hands: the closing regions of this call hold more than 2^32 point ids; split the sample list";
      return AG2_E_CAPACITY;
    }
    AG2_CUDA(c, cudaStreamSynchronize(c->stream));
    AG2_CUDA(c, c->box_pool.reserve((size_t)need + 1024));
    return kHandsRetry;
  }
  c->max_in_box = res[3];
  c->st.num_frames = (int64_t)c->h_pin[20];
  c->st.sum_neighbors_hands = (int64_t)stat[0];
  c->st.sum_cropped = (int64_t)stat[1];
  c->st.sum_in_box = (int64_t)stat[2];
  return AG2_OK;
}

int hands_compute(ag2_ctx* c) {
  for (int attempt = 0; attempt < 8; attempt++) {
    AG2_TRY(hands_launch(c));
    AG2_TRY(sync_and_finish(c));
    const int rc = hands_finish(c);
    if (rc != kHandsRetry) return rc;
  }
  return AG2_E_CAPACITY;
}

}  // namespace ag2
