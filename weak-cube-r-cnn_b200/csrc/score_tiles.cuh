// score_tiles.cuh -- warp-autonomous, culled form of the fused scoring kernel
// (included by score.cu after ScoreArgs).
//
// Why. A point can only lie inside a cube if its pixel lies inside the cube's
// projected box (perspective projection keeps convexity in front of the camera)
// and its depth lies between the cube's nearest and farthest corner; on the C2
// workload 55 % of the (proposal, ROI pixel) pairs fail one of the two. The
// thread-per-proposal layout (score_items_kernel) cannot use that: the 128
// proposals of a warp cover the whole ROI between them. Here the roles are
// swapped: a WARP owns a tile of depth pixels (lane = image row, register =
// column, 32 x 32 points), proposals stream past it in rounds of 32 -- culled
// lane-parallel against the tile's rows, columns and depth range, survivors then
// taken one at a time as warp-uniform parameters -- and inside a surviving tile
// the 8-register column groups the projected box cannot reach are skipped by
// uniform branches.
//
// The point test keeps the row-factored definition of score.cu bit for bit:
//     q_k = fma(z, fma(A_k, a_u, G_k), -O_k),  G_k = fma(B_k, b_v, C_k)
// G_k is now per (proposal, lane) -- the lane IS the row -- and a_u per register.
// Culling is conservative (box grown by one pixel, depth range by 1 mm per metre,
// both disabled when a corner is closer than 1 cm to the camera plane), so counts
// are identical to the unculled evaluation; tests compare them with the oracle,
// which never culls, and at full size with score_items_kernel.
//
// Work items, pulled by WARPS from two atomic queues (no CTA-level barrier anywhere):
//   tile item (box, chunk, tile, 128..1024 proposals of the chunk) -> point counts
//   seg item  (box, chunk, 32 proposals)                       -> hull fill pixel counts
// The warp that retires the last item of a (box, chunk) runs its epilogue (corners ->
// projection -> IoU -> score -> top-k) on its own; the warp that retires a box's last
// chunk merges the chunk candidates.
#pragma once

namespace wcr {

// proposals per tile item: ScoreArgs::ps = 1024 (the whole chunk) for shards with enough tiles to
// keep every warp busy, down to 128 for small ones (wcr_score_topk decides; 512 -> 1024 gave
// 11.14 -> 10.95 ms on C2)
constexpr int TL_SEGS = SC_CHUNK / 32;   // seg items per chunk
constexpr int TL_SCORE_OFF = 1024 + 3 * 32 * 4;   // per-warp scratch floats: a_u tile (also hull scratch) + survivor slots ...
constexpr int TL_WBUF = TL_SCORE_OFF + SC_CHUNK;  // ... + the epilogue's scores (hulls are filled while the scores are live)


// Tiling of a box's (sub-sampled) ROI: a tile is LR = 2^lrlog rows x (32/LR)*32
// columns; lane = (row in tile) + LR * phase, register r = column r*(32/LR) + phase.
struct TileGeom { int u0, v0, ncols, nrows, lrlog, n_rt, n_ct, n_tiles; };

__device__ __forceinline__ TileGeom tile_geom(const WcrBox &bx, int H, int W, int s, bool use_pts, int grp)
{
    const Roi r = box_roi(bx, H, W, s, use_pts);
    TileGeom g;
    g.u0 = r.u0; g.v0 = r.v0; g.ncols = r.ncols; g.nrows = r.nrows;
    int best = 4;
    long long best_cost = -1;
    const int order[3] = {4, 5, 3};   // ties go to the balanced 16-row tile
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int l = order[i], LR = 1 << l, gw = grp * (32 >> l);
        const long long cost = (long long)((g.nrows + LR - 1) / LR) * LR * (((g.ncols + gw - 1) / gw) * gw);
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = l; }
    }
    g.lrlog = best;
    g.n_rt = (g.nrows + (1 << best) - 1) >> best;
    const int tw = (32 >> best) * 32;
    g.n_ct = (g.ncols + tw - 1) / tw;
    g.n_tiles = g.nrows > 0 ? g.n_rt * g.n_ct : 0;
    return g;
}

// queue items of one (box, chunk); at least one so that somebody runs the epilogue
__device__ __forceinline__ int tile_items_per_chunk(const TileGeom &g, bool use_seg, int subs)
{
    return max(g.n_tiles * subs, 1) + (use_seg ? TL_SEGS : 0);
}

// Per-proposal constants of the point test, 48 B; the cull boxes live in their own
// int4 array so that a warp reads 32 of them in one coalesced request.
struct __align__(16) PParam {
    float4 a;   // A_0 A_1 A_2 -O_0   (rows scaled by 2^11 / half_k, see inside_sat)
    float4 b;   // B_0 B_1 B_2 -O_1
    float4 c;   // C_0 C_1 C_2 -O_2
};
// cull record (int4): ROI column range cl | ch << 16 and row range rl | rh << 16 the projected
// box can reach (inclusive; cl > ch: the proposal holds no point of this ROI), then the depth
// range [zlo, zhi] (float bits) of the cube's corners

// Depth range and number of the valid pixels of every cell of a 16 x 16 grid over a box's
// (sub-sampled) ROI: zgrid [n_boxes][256] (min, max), an empty cell holds (+inf, -inf); zcnt
// [n_boxes][256]; zblk [n_boxes][64] the ranges of the 2 x 2 blocks of cells; cellab [n_boxes][64]
// the pixel-ray slopes (the a_u, b_v of the point test) of the first and last column / row of every
// cell column / row: a_lo[16] a_hi[16] b_lo[16] b_hi[16]. (zcnt, zblk, cellab: only when asked for.)
// grid (n_boxes, 8) x 256 threads: a CTA takes cell rows 2y and 2y + 1; a thread walks its COLUMN
// down a cell row (coalesced across the CTA; valid depths are positive, so their bit patterns
// order like the values) and merges into the cell's shared-memory slot once per cell row.
// (The first form -- a warp per cell, lanes striding over the cell's pixels -- took 0.17 ms on C2.)
__global__ void __launch_bounds__(256)
score_zgrid_kernel(const WcrImage *__restrict__ images, const WcrBox *__restrict__ boxes, int stride,
                   int use_pts, float2 *__restrict__ zgrid, int *__restrict__ zcnt, float *__restrict__ cellab,
                   float2 *__restrict__ zblk)
{
    __shared__ unsigned s_mn[32], s_mx[32];
    __shared__ int s_n[32];
    const int b = blockIdx.x, y = blockIdx.y, tid = threadIdx.x;
    const WcrBox bx = boxes[b];
    const WcrImage im = images[bx.image];
    const Roi roi = box_roi(bx, im.H, im.W, stride, use_pts != 0);
    const int cw = max(1, (roi.ncols + 15) >> 4), chh = max(1, (roi.nrows + 15) >> 4);
    if (tid < 32) { s_mn[tid] = 0x7f800000u; s_mx[tid] = 0u; s_n[tid] = 0; }
    __syncthreads();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int cy = 2 * y + h, r0 = cy * chh, r1 = min(r0 + chh, roi.nrows);
        for (int c = tid; c < roi.ncols && r0 < r1; c += 256) {
            const float *col = im.depth + (size_t)(roi.v0 + r0 * stride) * im.depth_pitch + roi.u0 + c * stride;
            unsigned mn = 0x7f800000u, mx = 0u;
            int nv = 0;
            for (int r = r0; r < r1; ++r, col += (size_t)stride * im.depth_pitch) {
                const float d = __ldg(col);
                if (d > 0.f && d <= 3.402823466e38f) {         // the validity rule of tile_item
                    const unsigned u = __float_as_uint(d);
                    mn = min(mn, u); mx = max(mx, u); ++nv;
                }
            }
            if (nv) {
                const int slot = h * 16 + min(15, c / cw);
                atomicMin(&s_mn[slot], mn); atomicMax(&s_mx[slot], mx); atomicAdd(&s_n[slot], nv);
            }
        }
    }
    __syncthreads();
    if (tid < 32) {
        const int cell = y * 32 + tid;                        // rows 2y (slots 0..15) and 2y + 1 (16..31)
        const bool any = s_n[tid] > 0;
        zgrid[(size_t)b * 256 + cell] = make_float2(any ? __uint_as_float(s_mn[tid]) : INFINITY,
                                                    any ? __uint_as_float(s_mx[tid]) : -INFINITY);
        if (zcnt) zcnt[(size_t)b * 256 + cell] = s_n[tid];
    }
    if (zblk && tid < 8) {
        const int i = tid;
        unsigned mn = min(min(s_mn[2 * i], s_mn[2 * i + 1]), min(s_mn[16 + 2 * i], s_mn[16 + 2 * i + 1]));
        unsigned mx = max(max(s_mx[2 * i], s_mx[2 * i + 1]), max(s_mx[16 + 2 * i], s_mx[16 + 2 * i + 1]));
        const bool any = (s_n[2 * i] + s_n[2 * i + 1] + s_n[16 + 2 * i] + s_n[16 + 2 * i + 1]) > 0;
        zblk[(size_t)b * 64 + y * 8 + i] = make_float2(any ? __uint_as_float(mn) : INFINITY, any ? __uint_as_float(mx) : -INFINITY);
    }
    if (cellab && y == 0 && tid >= 32 && tid < 48) {
        const int i = tid - 32;
        const int cc0 = min(i * cw, max(roi.ncols - 1, 0)), cc1 = max(min(i * cw + cw, roi.ncols) - 1, cc0);
        const int rr0 = min(i * chh, max(roi.nrows - 1, 0)), rr1 = max(min(i * chh + chh, roi.nrows) - 1, rr0);
        float *ab = cellab + (size_t)b * 64;
        ab[i] = ((float)(roi.u0 + cc0 * stride) - im.cx) / im.fx;
        ab[16 + i] = ((float)(roi.u0 + cc1 * stride) - im.cx) / im.fx;
        ab[32 + i] = ((float)(roi.v0 + rr0 * stride) - im.cy) / im.fy;
        ab[48 + i] = ((float)(roi.v0 + rr1 * stride) - im.cy) / im.fy;
    }
}

// Mask pixels per cell of a 32 x 32 grid over a box's mask ROI, as a 33 x 33 integral table
// (mint[b][cy * 33 + cx] = mask pixels in cell rows < cy and cell columns < cx): the prep kernel
// bounds |mask & hull| by the mask pixels of the cells the hull's bounding box touches -- four loads
// per proposal. One CTA per box, one thread per cell (popcounts of the cell's bit ranges), then two
// prefix passes in shared memory.
constexpr int MG = 32;
__global__ void __launch_bounds__(MG * MG)
score_maskgrid_kernel(const WcrImage *__restrict__ images, const WcrBox *__restrict__ boxes,
                      const uint32_t *__restrict__ seg_bits, const int64_t *__restrict__ seg_off,
                      int *__restrict__ mint)
{
    __shared__ int s_c[MG][MG + 1];
    const int b = blockIdx.x, cx = threadIdx.x & (MG - 1), cy = threadIdx.x / MG;
    const WcrBox bx = boxes[b];
    const WcrImage im = images[bx.image];
    const Roi r = box_roi(bx, im.H, im.W, 1, true);
    SegBox sb;
    sb.u0 = r.u0; sb.v0 = r.v0; sb.ncols = r.ncols; sb.nrows = r.ncols > 0 ? r.nrows : 0;
    sb.wpr = (sb.ncols + 31) >> 5; sb.bits = seg_bits + seg_off[b]; sb.area = 0;
    const int gw = max(1, (sb.ncols + MG - 1) / MG), gh = max(1, (sb.nrows + MG - 1) / MG);
    const int c0 = cx * gw, c1 = min(c0 + gw, sb.ncols) - 1, r0 = cy * gh, r1 = min(r0 + gh, sb.nrows);
    int n = 0;
    if (c0 <= c1)
        for (int row = r0; row < r1; ++row) n += seg_popc_range(sb, row, c0, c1);
    s_c[cy][cx] = n;
    __syncthreads();
    if (cy == 0) {          // thread cx: running sum down column cx
        int a_ = 0;
        for (int i = 0; i < MG; ++i) { a_ += s_c[i][cx]; s_c[i][cx] = a_; }
    }
    __syncthreads();
    if (cx == 0) {          // thread cy: running sum along row cy
        int a_ = 0;
        for (int i = 0; i < MG; ++i) { a_ += s_c[cy][i]; s_c[cy][i] = a_; }
    }
    __syncthreads();
    int *out = mint + (size_t)b * (MG + 1) * (MG + 1);
    out[(cy + 1) * (MG + 1) + cx + 1] = s_c[cy][cx];
    if (cy == 0) out[cx + 1] = 0;
    if (cx == 0) out[(cy + 1) * (MG + 1)] = 0;
    if (threadIdx.x == 0) out[0] = 0;
}

// One thread per proposal: test constants (same expressions as score_items_kernel's
// prologue) and the conservative cull box.
__global__ void __launch_bounds__(256)
score_prep_kernel(const WcrImage *__restrict__ images, const WcrBox *__restrict__ boxes, int n_boxes,
                  const CubeSrc src, int N, int stride, int use_pts,
                  PParam *__restrict__ out, int4 *__restrict__ out_box,
                  const float2 *__restrict__ zgrid, uint8_t *__restrict__ out_mayhit,
                  float *__restrict__ out_siou, const int *__restrict__ zcnt, const float *__restrict__ cellab,
                  int *__restrict__ out_ubcnt, float *__restrict__ out_key, const float2 *__restrict__ zblk,
                  int2 *__restrict__ out_segb, unsigned *__restrict__ out_segwin, int e_iou, int e_pts)
{
    // (threads past the last proposal recompute proposal N - 1 and store nothing: the bound section
    // below is warp-cooperative)
    const int j_raw = blockIdx.y * 256 + threadIdx.x, b = blockIdx.x;
    const bool store = j_raw < N;
    const int j = min(j_raw, N - 1);
    const WcrBox bx = boxes[b];
    const WcrImage im = images[bx.image];
    const Roi roi = box_roi(bx, im.H, im.W, stride, use_pts != 0);
    float cube[15];
    if (src.cubes) {
        const float *cp = src.cubes + ((size_t)b * N + j) * 15;
#pragma unroll
        for (int i = 0; i < 15; ++i) cube[i] = ld_stream(cp + i);
    } else {   // sampled by the library: straight from the counter RNG, never stored
        const BoxCtx ctx = make_box_ctx(im, bx, src.priors, src.normals, src.z0[b], src.sp);
        sample_cube(ctx, nullptr, N, src.seed, (uint32_t)b, (uint32_t)j, cube);
    }
    // the IoU term of the score (corners -> K -> clamp -> min / max -> pairwise_iou), here where
    // every proposal has a thread of its own: the chunk epilogues (ONE warp per 1024 proposals)
    // then only combine it with the point count
    float siou_j = 1.f;
    if (out_siou) {
        const Intr Kc = {im.fx, im.fy, im.cx, im.cy};
        float lo0, hi0, lo1, hi1;
        clamp_bounds(bx.clamp0, lo0, hi0);
        clamp_bounds(bx.clamp1, lo1, hi1);
        float cxy[16];
        const ProjBox pbx = project_cube(cube, Kc, lo0, hi0, lo1, hi1, out_segb ? cxy : nullptr);
        siou_j = iou_2d(bx.x1, bx.y1, bx.x2, bx.y2, pbx);
        if (store) out_siou[(size_t)b * N + j] = siou_j;
        if (out_segb) {
            // Bounds for the mask term (seg_ratio_bound): the hull K that seg_group fills is the
            // convex hull of the corners clamped a second time (roi_heads.py:950-952), some of them
            // bumped by up to +7 pixels in x and y when they coincide (jarvis_march's duplicate
            // rule) -- bumps add points, the hull only grows.
            //   x = pixels of (K's bounding box, grown like seg_counts_warp's scan window and by
            //       the bumps) inside the mask's ROI: an upper bound of |mask & hull|;
            //   y = a lower bound of |hull|: K contains every face of the cube, so its area is at
            //       least the largest |signed area| of the six clamped face quadrilaterals; a
            //       convex region of area A and perimeter P holds at least A - 0.72 P pixel
            //       centres deeper than 0.01 px inside (unit squares around lattice points tile
            //       the plane; erode K by sqrt(2)/2 + 0.01), which is far more than the rounding
            //       of the fill's edge tests (~1e-3 px), and P <= the bounding box's perimeter.
            const float hx2 = (float)(int)(bx.clamp0 - 1.f), hy2 = (float)(int)(bx.clamp1 - 1.f);
            float sx[8], sy[8];
            bool fin = true;
            float mnx = INFINITY, mxx = -INFINITY, mny = INFINITY, mxy = -INFINITY;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                fin = fin && fabsf(cxy[2 * i]) <= 3.402823466e38f && fabsf(cxy[2 * i + 1]) <= 3.402823466e38f;
                sx[i] = fminf(fmaxf(cxy[2 * i], 0.f), hx2);
                sy[i] = fminf(fmaxf(cxy[2 * i + 1], 0.f), hy2);
                mnx = fminf(mnx, sx[i]); mxx = fmaxf(mxx, sx[i]);
                mny = fminf(mny, sy[i]); mxy = fmaxf(mxy, sy[i]);
            }
            int2 sb2 = make_int2(0x7fffffff, 0);
            unsigned swin = 0u | (32u << 6) | (0u << 12) | (32u << 18);   // whole grid: no cut
            if (fin && mxx < 1.0e6f && mxy < 1.0e6f) {
                const Roi r1 = box_roi(bx, im.H, im.W, 1, true);
                const int nr1 = r1.ncols > 0 ? r1.nrows : 0;
                const int x0 = max((int)floorf(mnx) - 1, r1.u0), x1 = min((int)ceilf(mxx) + 9, r1.u0 + r1.ncols - 1);
                const int y0 = max((int)floorf(mny) - 1, r1.v0), y1 = min((int)ceilf(mxy) + 9, r1.v0 + nr1 - 1);
                sb2.x = (x1 >= x0 && y1 >= y0) ? (x1 - x0 + 1) * (y1 - y0 + 1) : 0;
                // ... and, in the epilogues (the masks may still be on their way here), by the mask
                // pixels of the cells of a 32 x 32 grid that window touches (score_maskgrid_kernel):
                // the window's cell range, 6 bits each
                if (sb2.x > 0) {
                    const int gw = max(1, (r1.ncols + MG - 1) / MG), gh = max(1, (nr1 + MG - 1) / MG);
                    const int gx0 = min(MG - 1, (x0 - r1.u0) / gw), gx1 = min(MG - 1, (x1 - r1.u0) / gw) + 1;
                    const int gy0 = min(MG - 1, (y0 - r1.v0) / gh), gy1 = min(MG - 1, (y1 - r1.v0) / gh) + 1;
                    swin = (unsigned)gx0 | ((unsigned)gx1 << 6) | ((unsigned)gy0 << 12) | ((unsigned)gy1 << 18);
                }
                {
                    const int F[6][4] = {{0, 1, 2, 3}, {4, 5, 6, 7}, {0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {3, 0, 4, 7}};
                    float amax = 0.f;
#pragma unroll
                    for (int f = 0; f < 6; ++f) {
                        float acc = 0.f;
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int i0 = F[f][e], i1 = F[f][(e + 1) & 3];
                            acc += (sx[i0] - mnx) * (sy[i1] - mny) - (sx[i1] - mnx) * (sy[i0] - mny);
                        }
                        amax = fmaxf(amax, fabsf(acc));
                    }
                    // the fill counts no pixel outside the image: the reference clamps x with
                    // clamp0 and y with clamp1 (spaces.py:241-242; (H, W) in its call), so K may
                    // reach past the last row or column -- at most the part of its bounding box there
                    const float wbb = mxx - mnx + 1.f, hbb = mxy - mny + 1.f;
                    const float outside = wbb * fmaxf(mxy - (float)(im.H - 1), 0.f) + hbb * fmaxf(mxx - (float)(im.W - 1), 0.f);
                    const float lowf = 0.5f * amax * 0.999f - outside * 1.001f - 2.f * ((wbb + 16.f) + (hbb + 16.f)) - 16.f;
                    sb2.y = lowf > 0.f ? (int)lowf : 0;
                }
            }
            if (store) { out_segb[(size_t)b * N + j] = sb2; out_segwin[(size_t)b * N + j] = swin; }
        }
    }
    if (!use_pts) return;     // launched for the IoU term only
    PParam p;
    const bool live = roi.nrows > 0 && cube[3] > 0.f && cube[4] > 0.f && cube[5] > 0.f;
    float A[3], B[3], C[3], NO[3];
    if (live) {
        const float half[3] = {cube[5] / 2.f, cube[4] / 2.f, cube[3] / 2.f};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            A[k] = cube[6 + k] / half[k];
            B[k] = cube[9 + k] / half[k];
            C[k] = cube[12 + k] / half[k];
            NO[k] = -fmaf(C[k], cube[2], fmaf(B[k], cube[1], A[k] * cube[0]));
            A[k] *= 2048.f; B[k] *= 2048.f; C[k] *= 2048.f; NO[k] *= 2048.f;
        }
    } else {
#pragma unroll
        for (int k = 0; k < 3; ++k) { A[k] = B[k] = C[k] = 0.f; NO[k] = nanf_(); }
    }
    p.a = make_float4(A[0], A[1], A[2], NO[0]);
    p.b = make_float4(B[0], B[1], B[2], NO[1]);
    p.c = make_float4(C[0], C[1], C[2], NO[2]);
    int cl = 1, ch = 0, rl = 1, rh = 0;   // dead until proven reachable
    float zlo = -INFINITY, zhi = INFINITY; // depth range a point of the cube can have
    bool ortho = true;
    if (live) {
        // unclamped projection of the 8 corners; min/max and the nearest corner depth
        const float hx = cube[5] / 2.f, hy = cube[4] / 2.f, hz = cube[3] / 2.f;
        const float *R = cube + 6;
        float mnx = INFINITY, mxx = -INFINITY, mny = INFINITY, mxy = -INFINITY, mnz = INFINITY, mxz = -INFINITY;
        bool finite = true;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float sx, sy, sz;
            corner_signs(i, sx, sy, sz);
            const float lx = sx * hx, ly = sy * hy, lz = sz * hz;
            const float X = fmaf(R[2], lz, fmaf(R[1], ly, R[0] * lx)) + cube[0];
            const float Y = fmaf(R[5], lz, fmaf(R[4], ly, R[3] * lx)) + cube[1];
            const float Z = fmaf(R[8], lz, fmaf(R[7], ly, R[6] * lx)) + cube[2];
            const float px = fmaf(im.cx, Z, im.fx * X) / Z, py = fmaf(im.cy, Z, im.fy * Y) / Z;
            finite = finite && fabsf(px) <= 3.402823466e38f && fabsf(py) <= 3.402823466e38f &&
                     fabsf(Z) <= 3.402823466e38f;
            mnx = fminf(mnx, px); mxx = fmaxf(mxx, px);
            mny = fminf(mny, py); mxy = fmaxf(mxy, py);
            mnz = fminf(mnz, Z); mxz = fmaxf(mxz, Z);
        }
        cl = 0; ch = roi.ncols - 1; rl = 0; rh = roi.nrows - 1;   // no culling
        // The cull box and depth range come from the corners c + R (+-half); the point test accepts
        // {|R^T (p - c)|_k <= half_k}. The two describe the same solid only for an orthonormal R
        // (callers may hand in raw head poses): any drift beyond rounding switches culling off.
        float dev = 0.f;
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int c2 = a; c2 < 3; ++c2) {
                const float d = fmaf(R[6 + a], R[6 + c2], fmaf(R[3 + a], R[3 + c2], R[a] * R[c2]));
                dev = fmaxf(dev, fabsf(d - (a == c2 ? 1.f : 0.f)));
            }
        ortho = dev <= 1.0e-4f;      // NaN -> false
        if (finite && ortho && mnz >= 0.01f) {
            // pixel u = u0 + c*stride is reachable iff floor(mnx) - 1 <= u <= ceil(mxx) + 1
            const float big = 1.0e6f;
            const int x_lo = (int)floorf(fminf(fmaxf(mnx, -big), big)) - 1 - roi.u0;
            const int x_hi = (int)ceilf(fminf(fmaxf(mxx, -big), big)) + 1 - roi.u0;
            const int y_lo = (int)floorf(fminf(fmaxf(mny, -big), big)) - 1 - roi.v0;
            const int y_hi = (int)ceilf(fminf(fmaxf(mxy, -big), big)) + 1 - roi.v0;
            // c >= ceil(x_lo / s), c <= floor(x_hi / s)   (floor division for negatives)
            const int s = stride;
            cl = max(cl, x_lo <= 0 ? 0 : (x_lo + s - 1) / s);
            ch = min(ch, x_hi < 0 ? -1 : x_hi / s);
            rl = max(rl, y_lo <= 0 ? 0 : (y_lo + s - 1) / s);
            rh = min(rh, y_hi < 0 ? -1 : y_hi / s);
            // a point inside the cube has its depth between the nearest and the farthest corner;
            // 1 mm per metre of slack covers the rounding of the test and of the corners
            const float slack = 1.0e-3f * (fabsf(mnz) + fabsf(mxz)) + 1.0e-4f;
            zlo = mnz - slack; zhi = mxz + slack;
        }
    }
    // Can the cube hold any valid pixel at all? Its corner depth range against the depth ranges of
    // the ROI cells (16 x 16 grid, score_zgrid_kernel) its cull box overlaps. "No" is a proof that
    // the point count is 0: such a proposal is dead for the tile items and its mask term is lazy.
    bool mayhit = cl <= ch && rl <= rh;
    // Point term on demand: an upper bound of the inside-count = the valid points of the cells the
    // cube can reach (a point inside the cube lies inside the cull box and the depth range; the cell
    // that holds it then overlaps both). Cubes the cull does not apply to can hold every point.
    int ubc = mayhit ? 0x7fffffff : 0;
    if (mayhit && zgrid && ortho && zcnt) {
        // 2 x 2 blocks of cells that overlap the cull box and the depth range are tested against the
        // cube itself: interval arithmetic on the very function the point test evaluates,
        //     q_k = z (A_k a_u + B_k b_v + C_k) - O_k,   inside iff |q_k| <= 2048 for k = 0, 1, 2,
        // over the block's pixel rectangle and the depths both the block and the cube can hold (the
        // points of a block lie on a surface patch: its narrow depth range is what makes the test
        // sharp). If some q_k stays outside [-2048, 2048] over the whole block -- with a margin for
        // the rounding of both evaluations -- no point of the block is inside. (Valid for any R: it
        // bounds the test itself.) Blocks rather than cells: 29 instead of 108 tests per proposal on
        // C2 for 70 % of the pruning the per-cell test adds (tools/prune_study.py).
        const int cw = max(1, (roi.ncols + 15) >> 4), chh = max(1, (roi.nrows + 15) >> 4);
        const float2 *zg = zgrid + (size_t)b * 256;
        const float2 *zb = zblk + (size_t)b * 64;
        const int *zc = zcnt + (size_t)b * 256;
        const int cx0 = cl / cw, cx1 = min(15, ch / cw), cy0 = rl / chh, cy1 = min(15, rh / chh);
        const float Ak[3] = {p.a.x, p.a.y, p.a.z}, Bk[3] = {p.b.x, p.b.y, p.b.z}, Ck[3] = {p.c.x, p.c.y, p.c.z};
        const float Nk[3] = {p.a.w, p.b.w, p.c.w};
        const float *ab = cellab + (size_t)b * 64;   // a_lo[16] a_hi[16] b_lo[16] b_hi[16] (score_zgrid_kernel)
        const float a_abs = fmaxf(fabsf(__ldg(ab)), fabsf(__ldg(ab + 31)));
        const float b_abs = fmaxf(fabsf(__ldg(ab + 32)), fabsf(__ldg(ab + 63)));
        float th[3];   // one margin per axis: 2.5 absolute + 1e-5 of the largest magnitude that can appear
#pragma unroll
        for (int k = 0; k < 3; ++k)
            th[k] = 2050.5f + 1.0e-5f * (fabsf(zhi) * (fabsf(Ak[k]) * a_abs + fabsf(Bk[k]) * b_abs + fabsf(Ck[k])) + fabsf(Nk[k]));
        ubc = 0;
        for (int by = cy0 >> 1; by <= (cy1 >> 1); ++by) {
            const float b_lo = __ldg(ab + 32 + 2 * by), b_hi = __ldg(ab + 48 + 2 * by + 1);
            float gl[3], gh[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float t1 = Bk[k] * b_lo, t2 = Bk[k] * b_hi;
                gl[k] = fminf(t1, t2) + Ck[k]; gh[k] = fmaxf(t1, t2) + Ck[k];
            }
            for (int bxi = cx0 >> 1; bxi <= (cx1 >> 1); ++bxi) {
                const float2 c = __ldg(zb + by * 8 + bxi);
                if (zhi < c.x || zlo > c.y) continue;              // (an empty block holds (+inf, -inf))
                const float z0c = fmaxf(c.x, zlo), z1c = fminf(c.y, zhi);
                const float a_lo = __ldg(ab + 2 * bxi), a_hi = __ldg(ab + 16 + 2 * bxi + 1);
                bool hit = true;
#ifndef WCR_EXP_NO_SLAB
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float t1 = Ak[k] * a_lo, t2 = Ak[k] * a_hi;
                    const float g_lo = fminf(t1, t2) + gl[k], g_hi = fmaxf(t1, t2) + gh[k];
                    const float q_lo = fminf(z0c * g_lo, z1c * g_lo) + Nk[k], q_hi = fmaxf(z0c * g_hi, z1c * g_hi) + Nk[k];
                    hit = hit && !(q_lo > th[k] || q_hi < -th[k]);
                }
#endif
                if (!hit) continue;
                // the block's cells inside the cull box whose depth range meets the cube's
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int cy = 2 * by + (q >> 1), cx = 2 * bxi + (q & 1);
                    if (cy < cy0 || cy > cy1 || cx < cx0 || cx > cx1) continue;
                    const float2 cc = __ldg(zg + cy * 16 + cx);
                    if (!(zhi < cc.x || zlo > cc.y)) ubc += __ldg(zc + cy * 16 + cx);
                }
            }
        }
        mayhit = ubc > 0;   // an empty cell holds (+inf, -inf): same verdict as the scan below
    } else if (mayhit && zgrid && ortho) {
        const int cw = max(1, (roi.ncols + 15) >> 4), chh = max(1, (roi.nrows + 15) >> 4);
        const float2 *zg = zgrid + (size_t)b * 256;
        mayhit = false;
        const int cx0 = cl / cw, cx1 = min(15, ch / cw);
        for (int cy = rl / chh; cy <= min(15, rh / chh) && !mayhit; ++cy) {
            // a whole row of cells per trip, no exit inside it: the loads are independent and overlap
#pragma unroll 4
            for (int cx = cx0; cx <= cx1; ++cx) {
                const float2 c = __ldg(zg + cy * 16 + cx);
                mayhit = mayhit || !(zhi < c.x || zlo > c.y);
            }
        }
    }
    if (!store) return;
    if (out_mayhit) out_mayhit[(size_t)b * N + j] = mayhit ? 1 : 0;
    if (zcnt) {
        out_ubcnt[(size_t)b * N + j] = ubc;
        // ordering key of the first pass (the score bound up to the common factor 1 / P_b^e_pts);
        // only the ORDER it induces is used, the pruning test recomputes the bound exactly
        out_key[(size_t)b * N + j] = ipow(siou_j, e_iou) * ipow((float)min(ubc, 1 << 24), e_pts);
    }
    if (!mayhit) { cl = 0xffff; ch = 0; rl = 0xffff; rh = 0; }   // fails every overlap test
    out[(size_t)b * N + j] = p;
    // 16 bytes per proposal: ROI column range, row range (16 bits each, inclusive) and depth range
    out_box[(size_t)b * N + j] = make_int4(cl | (ch << 16), rl | (rh << 16), __float_as_int(zlo), __float_as_int(zhi));
}

// Point term on demand: the proposals of a (box, chunk) by descending key (ties: lower index
// first; NaN keys first), one CTA of 1024 threads per chunk, bitonic sort of 64-bit composites
// (ordered key bits << 16 | 1023 - index) in shared memory. Slots past the chunk's end sort last.
__global__ void __launch_bounds__(1024)
score_order_kernel(const float *__restrict__ key, int N, int n_chunks, uint16_t *__restrict__ order)
{
    __shared__ unsigned long long s_v[SC_CHUNK];
    const int bc = blockIdx.x, b = bc / n_chunks, chunk = bc - b * n_chunks, t = threadIdx.x;
    const int chunk_n = min(SC_CHUNK, N - chunk * SC_CHUNK);
    unsigned long long v = 0ull;
    if (t < chunk_n) {
        unsigned u = __float_as_uint(key[(size_t)b * N + (size_t)chunk * SC_CHUNK + t]);
        u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);        // monotone map float -> unsigned
        v = ((unsigned long long)u << 16) | (unsigned)(SC_CHUNK - 1 - t) | (1ull << 48);
    }
    s_v[t] = v;
    __syncthreads();
    for (int k = 2; k <= SC_CHUNK; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int p = t ^ j;
            if (p > t) {
                const unsigned long long x = s_v[t], y = s_v[p];
                const bool desc = (t & k) == 0;                 // descending blocks -> descending array
                if (desc ? x < y : x > y) { s_v[t] = y; s_v[p] = x; }
            }
            __syncthreads();
        }
    order[(size_t)bc * SC_CHUNK + t] = (uint16_t)(SC_CHUNK - 1 - (int)(s_v[t] & 0xffffu));
}

// One CTA: queue items per box -> exclusive prefix (tile path), in the ORDER the boxes are served:
// boxes with the most tiles first (up to 2048 boxes; beyond that, input order). The grid's last
// items then belong to small boxes -- short tile items, small hulls -- and the warps run dry within
// tens of microseconds of each other instead of waiting for one 2000-pixel-wide ROI claimed last.
// box_order[pos] = box served at position pos; prefix[pos] = first item of that box.
__global__ void __launch_bounds__(1024)
score_plan_tiles_kernel(const WcrImage *__restrict__ images, const WcrBox *__restrict__ boxes,
                        int n_boxes, int n_chunks, int stride, int use_pts, int use_seg, int grp, int subs,
                        int *__restrict__ prefix, int *__restrict__ box_order)
{
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    __shared__ unsigned long long s_key[2048];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const bool sorted = n_boxes <= 2048;
    int K = 32;                                    // slots sorted: the next power of two
    while (K < n_boxes) K <<= 1;
    if (sorted) {
        for (int i = tid; i < K; i += 1024) {
            unsigned long long v = 0ull;
            if (i < n_boxes) {
                const WcrBox bx = boxes[i];
                const WcrImage im = images[bx.image];
                const int n = tile_items_per_chunk(tile_geom(bx, im.H, im.W, stride, use_pts != 0, grp), use_seg != 0, subs);
                v = ((unsigned long long)(unsigned)n << 16) | (unsigned)(2047 - i) | (1ull << 48);   // ties: lower index first
            }
            s_key[i] = v;
        }
        __syncthreads();
        for (int k = 2; k <= K; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int t = tid; t < K; t += 1024) {
                    const int p = t ^ j;
                    if (p > t) {
                        const unsigned long long x = s_key[t], y = s_key[p];
                        if (((t & k) == 0) ? x < y : x > y) { s_key[t] = y; s_key[p] = x; }
                    }
                }
                __syncthreads();
            }
    }
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < n_boxes; base += 1024) {
        const int pos = base + tid;
        int n = 0;
        if (pos < n_boxes) {
            const int b = sorted ? 2047 - (int)(s_key[pos] & 0xffffu) : pos;
            box_order[pos] = b;
            const WcrBox bx = boxes[b];
            const WcrImage im = images[bx.image];
            n = tile_items_per_chunk(tile_geom(bx, im.H, im.W, stride, use_pts != 0, grp), use_seg != 0, subs) * n_chunks;
        }
        int incl = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            const int t = s_warp[lane];
            int w = t;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int x = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += x;
            }
            s_warp[lane] = w - t;
        }
        __syncthreads();
        const int carry = s_carry;
        if (pos < n_boxes) prefix[pos] = carry + s_warp[wid] + incl - n;
        __syncthreads();
        if (tid == 1023) s_carry = carry + s_warp[wid] + incl;
        __syncthreads();
    }
    if (tid == 0) prefix[n_boxes] = s_carry;
}

}  // namespace wcr
