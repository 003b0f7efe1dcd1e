// score_tiles_kernel.cuh -- the persistent warp-autonomous scoring kernel
// (see score_tiles.cuh for the design). Included by score.cu.
#pragma once

namespace wcr {

__device__ __forceinline__ float4 ldg_f4(const float4 *p) { return __ldg(p); }

// acc + (m > 2^11 or m is NaN ? 1 : 0) for m = max_k |q_k| >= 0 on the INTEGER pipe: non-negative
// floats order like their bit patterns (NaN above everything), so the sign bit of
// bits(2^11) - bits(m) is the "outside" indicator; IADD3 + LEA.HI, no FMA-pipe slot.
__device__ __forceinline__ int add_outside(int acc, float q0, float q1, float q2)
{
    float m;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(fabsf(q0)), "f"(fabsf(q1)), "f"(fabsf(q2)));
    int t;
    unsigned sgn;
    asm("sub.s32 %0, 0x45000000, %1;" : "=r"(t) : "r"(__float_as_int(m)));
    asm("shr.u32 %0, %1, 31;" : "=r"(sgn) : "r"(t));
    return acc + (int)sgn;
}

// ---- tile item: one warp, 32 x 32 depth pixels (z in registers, a_u in the warp's
// shared scratch), proposals stream past in rounds of 32:
//   lane = proposal: cull box against the tile -> ballot of survivors and of the column
//                    groups each one reaches; survivors' constants -> shared slots
//   lane = row:      survivors one at a time as warp-uniform parameters (broadcast LDS),
//                    ballots and bit scans keep the control flow in uniform registers
//   lane = proposal: one coalesced RED of the round's counts
template <int GRP, bool F2, bool ICNT>
__device__ __forceinline__ void tile_item(const ScoreArgs &a, int b, const WcrImage &im,
                                          const TileGeom &tg, int chunk, int tile, int sub,
                                          float *wbuf, int lane,
                                          const uint16_t *list = nullptr, int n_list = 0, bool count_valid = true)
{
    constexpr int NG = 32 / GRP;
    const int s = a.sc.point_stride;
    const int lrlog = tg.lrlog, LR = 1 << lrlog, nph = 32 >> lrlog;
    const int rt = tile / tg.n_ct, ct = tile - rt * tg.n_ct;
    const int tr0 = rt << lrlog, tc0 = ct * (nph * 32);
    const int my_row = tr0 + (lane & (LR - 1)), phase = lane >> lrlog;
    const bool row_ok = my_row < tg.nrows;
    const int v = tg.v0 + my_row * s;

    // the proposals of the item: slots sub * ps .. of the chunk, or (point term on demand) the
    // n_list chunk-relative indices of `list` (written during this launch: read through L2)
    const int chunk_n = min(SC_CHUNK, a.N - chunk * SC_CHUNK);
    const int n = list ? n_list : min(a.ps, chunk_n - sub * a.ps);
    if (n <= 0) return;
    const size_t first = (size_t)b * a.N + (size_t)chunk * SC_CHUNK + (list ? 0 : sub * a.ps);
    const PParam *__restrict__ pp = a.ws.params + first;
    const int4 *__restrict__ cbx = a.ws.cullbox + first;
    int loc_next = min(lane, n - 1);
    if (list) loc_next = (int)__ldcg(list + loc_next);
    int4 bb_next = __ldg(cbx + loc_next);

    const float *drow = im.depth + (size_t)(row_ok ? v : 0) * im.depth_pitch;
    float4 *av4 = reinterpret_cast<float4 *>(wbuf);          // [8][32]: a_u of registers 4*r4 .. 4*r4+3
    float4 *slot = reinterpret_cast<float4 *>(wbuf + 1024);  // [3][32]: constants of the round's survivors
    u64 z2[16];   // z of registers (2i, 2i+1) as one 64-bit pair
    int valid = 0;
#pragma unroll
    for (int r4 = 0; r4 < 8; ++r4) {
        float a4[4], z4[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int r = r4 * 4 + t;
            const int col = tc0 + r * nph + phase;
            const int u = tg.u0 + col * s;
            a4[t] = ((float)u - im.cx) / im.fx;
            float d = 0.f;
            if (row_ok && col < tg.ncols) d = __ldg(drow + u);
            const bool ok = d > 0.f && d <= 3.402823466e38f;
            z4[t] = ok ? d : nanf_();
            valid += ok;
        }
        z2[r4 * 2] = pack2(z4[0], z4[1]);
        z2[r4 * 2 + 1] = pack2(z4[2], z4[3]);
        av4[r4 * 32 + lane] = make_float4(a4[0], a4[1], a4[2], a4[3]);
    }
    const float bv = ((float)v - im.cy) / im.fy;
    // depth range of the tile's valid points (fminf / fmaxf drop the NaN of invalid ones)
    float t_zmin = INFINITY, t_zmax = -INFINITY;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        float lo, hi;
        unpack2(z2[i], lo, hi);
        t_zmin = fminf(t_zmin, fminf(lo, hi)); t_zmax = fmaxf(t_zmax, fmaxf(lo, hi));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        t_zmin = fminf(t_zmin, __shfl_xor_sync(0xffffffffu, t_zmin, o));
        t_zmax = fmaxf(t_zmax, __shfl_xor_sync(0xffffffffu, t_zmax, o));
    }
    if (sub == 0 && count_valid) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) valid += __shfl_xor_sync(0xffffffffu, valid, o);
        if (lane == 0 && valid) atomicAdd(a.ws.npts + (size_t)b * a.n_chunks + chunk, valid);
    }
    int *__restrict__ cnt = a.ws.count + first;
    int executed = 0;
    const int t_r1 = min(tr0 + LR, tg.nrows) - 1, t_c1 = min(tc0 + nph * 32, tg.ncols) - 1;
    const int gshift = (GRP == 32 ? 5 : GRP == 16 ? 4 : GRP == 8 ? 3 : 2) + (5 - lrlog);   // log2(GRP * nph)

#pragma unroll 1
    for (int base = 0; base < n; base += 32) {
        const int4 bb = bb_next;
        const int j = base + lane, jl = loc_next;   // slot in the item, index in the chunk
        loc_next = min(j + 32, n - 1);
        if (list) loc_next = (int)__ldcg(list + loc_next);
        bb_next = __ldg(cbx + loc_next);
        const int cl = bb.x & 0xffff, ch = (unsigned)bb.x >> 16, rl = bb.y & 0xffff, rh = (unsigned)bb.y >> 16;
        const float zlo = __int_as_float(bb.z), zhi = __int_as_float(bb.w);
        const bool keep = j < n && !(cl > t_c1 || ch < tc0 || rl > t_r1 || rh < tr0) &&
                          !(zhi < t_zmin || zlo > t_zmax);
        unsigned m = __ballot_sync(0xffffffffu, keep);
        if (m == 0) continue;
        const int g_lo = (max(cl, tc0) - tc0) >> gshift, g_hi = (min(ch, t_c1) - tc0) >> gshift;
        if (keep) executed += (g_hi - g_lo + 1) * GRP;   // registers this proposal is tested on
        unsigned mg[NG];
#pragma unroll
        for (int g = 0; g < NG; ++g) mg[g] = __ballot_sync(0xffffffffu, keep && g >= g_lo && g <= g_hi);
        __syncwarp();   // the previous round's slot reads are done
        if (keep) {
            const float4 pa = ldg_f4(&pp[jl].a), pb = ldg_f4(&pp[jl].b), pc = ldg_f4(&pp[jl].c);
            slot[lane] = pa; slot[32 + lane] = pb; slot[64 + lane] = pc;
        }
        __syncwarp();
        int mycnt = 0;
        while (m) {
            const int p = __ffs(m) - 1;
            m &= m - 1;
            const float4 ca = slot[p], cb4 = slot[32 + p], cc = slot[64 + p];
            const float G0 = fmaf(cb4.x, bv, cc.x), G1 = fmaf(cb4.y, bv, cc.y), G2 = fmaf(cb4.z, bv, cc.z);
            int c;
            if (F2) {
                // Blackwell packed FP32: registers (2i, 2i+1) of the tile are two points of the
                // same row, so one fma.rn.f32x2 advances both tests (same IEEE roundings).
                const u64 A0 = pack2(ca.x, ca.x), A1 = pack2(ca.y, ca.y), A2 = pack2(ca.z, ca.z);
                const u64 N0 = pack2(ca.w, ca.w), N1 = pack2(cb4.w, cb4.w), N2 = pack2(cc.w, cc.w);
                const u64 GG0 = pack2(G0, G0), GG1 = pack2(G1, G1), GG2 = pack2(G2, G2);
                // steps of 8 registers (two LDS.128 of a_u, the next step's issued before this
                // step's arithmetic; with 168 registers ptxas hoists the whole a_u tile out of
                // the survivor loop). ICNT: indicator and counter on the integer pipe ("outside"
                // bits are summed and taken from the number of registers visited); otherwise
                // FFMA.SAT indicator + FADD2 counter on the FMA pipe.
                int oc[4] = {0, 0, 0, 0}, visited = 0;
                u64 acc2[4] = {pack2(0.f, 0.f), pack2(0.f, 0.f), pack2(0.f, 0.f), pack2(0.f, 0.f)};
                constexpr int SPG = GRP >= 8 ? GRP / 8 : 1;   // steps per column group (the FFMA2 path needs GRP >= 8)
                const ulonglong2 *avq = reinterpret_cast<const ulonglong2 *>(av4) + lane;
                ulonglong2 n0 = avq[0], n1 = avq[32];
#pragma unroll
                for (int st = 0; st < 4; ++st) {
                    const ulonglong2 c0 = n0, c1 = n1;
                    if (st < 3) { n0 = avq[(2 * st + 2) * 32]; n1 = avq[(2 * st + 3) * 32]; }
                    if ((mg[st / SPG] >> p) & 1u) {
                        visited += 8;
                        const u64 avp[4] = {c0.x, c0.y, c1.x, c1.y};
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            const u64 zz = z2[st * 4 + t];
                            float q0a, q0b, q1a, q1b, q2a, q2b;
                            unpack2(ffma2(zz, ffma2(A0, avp[t], GG0), N0), q0a, q0b);
                            unpack2(ffma2(zz, ffma2(A1, avp[t], GG1), N1), q1a, q1b);
                            unpack2(ffma2(zz, ffma2(A2, avp[t], GG2), N2), q2a, q2b);
                            if (ICNT) {
                                oc[t] = add_outside(oc[t], q0a, q1a, q2a);
                                oc[t] = add_outside(oc[t], q0b, q1b, q2b);
                            } else {
                                acc2[t] = fadd2(acc2[t], pack2(inside_sat(q0a, q1a, q2a), inside_sat(q0b, q1b, q2b)));
                            }
                        }
                    }
                }
                if (ICNT) {
                    c = visited - ((oc[0] + oc[1]) + (oc[2] + oc[3]));
                } else {
                    acc2[0] = fadd2(acc2[0], acc2[2]);
                    acc2[1] = fadd2(acc2[1], acc2[3]);
                    float s0, s1, s2, s3;
                    unpack2(acc2[0], s0, s1);
                    unpack2(acc2[1], s2, s3);
                    c = (int)((s0 + s1) + (s2 + s3));
                }
            } else {
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int g = 0; g < NG; ++g) {
                    if ((mg[g] >> p) & 1u) {
#pragma unroll
                        for (int h = 0; h < GRP / 4; ++h) {
                            const float4 a4 = av4[(g * (GRP / 4) + h) * 32 + lane];
                            const float avv[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                float zlo, zhi;
                                unpack2(z2[(g * GRP + h * 4 + t) / 2], zlo, zhi);
                                const float zr = (t & 1) ? zhi : zlo;
                                const float q0 = fmaf(zr, fmaf(ca.x, avv[t], G0), ca.w);
                                const float q1 = fmaf(zr, fmaf(ca.y, avv[t], G1), cb4.w);
                                const float q2 = fmaf(zr, fmaf(ca.z, avv[t], G2), cc.w);
                                acc[t] += inside_sat(q0, q1, q2);
                            }
                        }
                    }
                }
                c = (int)((acc[0] + acc[1]) + (acc[2] + acc[3]));
            }
            c = __reduce_add_sync(0xffffffffu, c);
            if (lane == p) mycnt = c;
        }
        if (mycnt) atomicAdd(cnt + jl, mycnt);
    }
    // bookkeeping for the roofline report: lane-tests actually issued by this item
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) executed += __shfl_xor_sync(0xffffffffu, executed, o);
    if (lane == 0) atomicAdd(a.ws.executed, (unsigned long long)executed * 32ull);
    __syncwarp();
}

// The 15 floats of proposal j of box b: read from the caller's array, or -- when the library
// samples (cubes == NULL) -- REGENERATED from the counter RNG wherever they are needed (Philox is
// keyed by (seed; proposal, box, attribute, try): no state, so a proposal is never stored; the
// records carry the winners' cubes). Out of line: three call sites, none in the tile loop.
__device__ __noinline__ void fetch_cube(const CubeSrc src, const WcrImage im, const WcrBox bx, int b, int j,
                                        float *cube)
{
    if (src.cubes) {
        const float *p = src.cubes + ((size_t)b * src.N + j) * 15;
#pragma unroll
        for (int i = 0; i < 15; ++i) cube[i] = ld_stream(p + i);
        return;
    }
    const BoxCtx ctx = make_box_ctx(im, bx, src.priors, src.normals, src.z0[b], src.sp);
    float c[15];
    sample_cube(ctx, nullptr, src.N, src.seed, (uint32_t)b, (uint32_t)j, c);
#pragma unroll
    for (int i = 0; i < 15; ++i) cube[i] = c[i];
}

// ---- seg item: hull fill pixel counts of 32 proposals (lane = proposal for the march,
// then the warp walks the 32 hulls with lanes = pixel rows) --------------------------
// q = the lane's proposal within the chunk, or -1 for a lane without one.
// Out of line and fed by value: two call sites (seg items, the epilogue's lazy fill) share one
// copy of the hull code (measured faster than inlining it into the seg items: 10.76 vs 10.87 ms),
// and the kernel's parameter block never has to be copied to local memory.
__device__ __noinline__ void seg_group(const CubeSrc src, const uint32_t *__restrict__ seg_bits_b,
                                       int2 *__restrict__ segcnt, int b, const WcrBox bx, const WcrImage im,
                                       int chunk, int q, float *wbuf, int lane)
{
    const Intr K = {im.fx, im.fy, im.cx, im.cy};
    float lo0, hi0, lo1, hi1;
    clamp_bounds(bx.clamp0, lo0, hi0);
    clamp_bounds(bx.clamp1, lo1, hi1);
    SegBox sb;
    const Roi r1 = box_roi(bx, im.H, im.W, 1, true);
    sb.u0 = r1.u0; sb.v0 = r1.v0; sb.ncols = r1.ncols; sb.nrows = r1.ncols > 0 ? r1.nrows : 0;
    sb.wpr = (sb.ncols + 31) >> 5;
    sb.bits = seg_bits_b;
    sb.area = 0;
    float *scratch = wbuf + lane;
    const Strided spx = {scratch, 32}, spy = {scratch + 8 * 32, 32};
    const Strided shx = {scratch + 16 * 32, 32}, shy = {scratch + 24 * 32, 32};
    const float hx2 = (float)(int)(bx.clamp0 - 1.f), hy2 = (float)(int)(bx.clamp1 - 1.f);
    int m_own = -1;
    if (q >= 0) {
        float cube[15], cxy[16];
        fetch_cube(src, im, bx, b, chunk * SC_CHUNK + q, cube);
        project_cube(cube, K, lo0, hi0, lo1, hi1, cxy);
        bool finite = true;   // second clamp of the corners (roi_heads.py:950-952)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            finite = finite && fabsf(cxy[2 * i]) <= 3.402823466e38f && fabsf(cxy[2 * i + 1]) <= 3.402823466e38f;
            spx(i) = fminf(fmaxf(cxy[2 * i], 0.f), hx2);
            spy(i) = fminf(fmaxf(cxy[2 * i + 1], 0.f), hy2);
        }
        if (finite) {
            m_own = jarvis_march(spx, spy, 8, shx, shy);
            if (m_own > 0) hull_edges(shx, shy, m_own, spx, spy);   // corners are dead: reuse
        }
    }
    __syncwarp();
    int my_inter = 0, my_bube = 0;
#pragma unroll 1
    for (int p = 0; p < 32; ++p) {
        const int mp = __shfl_sync(0xffffffffu, m_own, p);
        if (mp < 1) continue;
        const Strided exp_ = {wbuf + p, 32}, eyp = {wbuf + p + 8 * 32, 32};
        const Strided hxp = {wbuf + p + 16 * 32, 32}, hyp = {wbuf + p + 24 * 32, 32};
        int it, bu;
        seg_counts_warp(hxp, hyp, exp_, eyp, mp, im.H, im.W, sb, it, bu);
        if (lane == p) { my_inter = it; my_bube = bu; }
    }
    __syncwarp();
    if (q >= 0) segcnt[(size_t)b * src.N + chunk * SC_CHUNK + q] = make_int2(my_inter, my_bube);
}

// seg item: proposals grp*32 .. grp*32+31 of the chunk (only when EVERY mask term is wanted:
// per-proposal outputs, or mod_mask_iou whose ratio is not bounded by 1; otherwise the epilogue
// evaluates the few it needs, see chunk_epilogue_warp).
__device__ __forceinline__ void seg_item(const ScoreArgs &a, int b, const WcrBox &bx, const WcrImage &im,
                                         int chunk, int grp, float *wbuf, int lane)
{
    const int chunk_n = min(SC_CHUNK, a.N - chunk * SC_CHUNK);
    if (grp * 32 >= chunk_n) return;
    int q = grp * 32 + lane;
    if (q >= chunk_n) q = -1;
    if (lane == 0) atomicAdd(a.ws.hulls, (unsigned long long)min(32, chunk_n - grp * 32));
    seg_group(a.src, a.seg_bits + a.seg_off[b], a.ws.segcnt, b, bx, im, chunk, q, wbuf, lane);
}

// Terms of one proposal (epilogue and record writer share it).
struct PropTerms { float score, siou, spts, sseg; int count; ProjBox pb; };

__device__ __forceinline__ PropTerms prop_terms(const ScoreArgs &a, const float *cube, size_t o,
                                                const WcrBox &bx, const Intr &K, float lo0, float hi0,
                                                float lo1, float hi1, int npts, int area,
                                                bool with_seg = true)
{
    PropTerms t;
    t.count = __ldcg(a.ws.count + o);
    t.pb = project_cube(cube, K, lo0, hi0, lo1, hi1, nullptr);
    t.siou = iou_2d(bx.x1, bx.y1, bx.x2, bx.y2, t.pb);
    t.spts = npts > 0 ? (float)t.count / (float)npts : 0.f;
    t.sseg = 1.f;
    if (a.sc.e_seg > 0 && with_seg) {
        const int2 sc2 = __ldcg(a.ws.segcnt + o);
        t.sseg = seg_ratio(sc2.x, sc2.y, area, a.sc.seg_mode);
    }
    const float mb = a.sc.map_beta;
    t.score = (ipow(shape_term(t.siou, mb), a.sc.e_iou) * ipow(shape_term(t.spts, mb), a.sc.e_pts)) *
              ipow(shape_term(t.sseg, mb), a.sc.e_seg);
    return t;
}

// ---- epilogue of a (box, chunk): terms -> score -> top-k -> records ----------------------
//
// Full evaluation (a.seg_lazy == 0): the chunk's seg items have filled every hull; the warp that
// retires the chunk's last item computes every term, takes the top-k and writes the records.
//
// Mask term on demand (a.seg_lazy: mask_iou mode, no per-proposal scores / ratios asked for).
// The score is ub * sseg^e with ub = iou^a * pts^b and 0 <= sseg <= 1, every factor rounded the
// same way whether sseg is known or not, so score <= ub in floating point too. Only proposals
// that can still reach the top-k need their hull filled:
//   phase A (the warp that retires the chunk's last tile item): ub of every proposal -> ws.ub;
//           the 32 best bounds are evaluated on the spot; T = k-th best ACTUAL score among them;
//           every other proposal with ub >= T and ub > 0 (a zero bound is a zero score whatever
//           the mask says) goes to the chunk's candidate list, which is cut into batches of 32
//           and PUBLISHED on a dynamic queue that every warp of the grid serves before it takes
//           a new tile item -- a chunk's hulls are filled by many warps, not by one;
//   batch   (any warp): hull fill of 32 candidates, actual score -> ws.ub;
//   phase B (the warp that retires the chunk's last batch): top-k over ws.ub. An unevaluated
//           proposal has score <= ub < T <= k-th best actual score: it cannot be in the top-k,
//           not even through a tie; unevaluated zero bounds ARE their exact score. So the top-k
//           over the stored values is the top-k of the full evaluation, bit for bit
//           (tests/test_score_gpu.py: on-demand == full evaluation == oracle).
// Winners with a zero score get their mask term filled in for the record's s_seg field.
struct ChunkCtx {
    Intr K;
    float lo0, hi0, lo1, hi1;
    int npts, area;
    size_t cbase;     // first proposal of the chunk in [n_boxes][N] arrays
    int chunk_n;      // proposals in the chunk
};

__device__ __forceinline__ ChunkCtx chunk_ctx(const ScoreArgs &a, int b, const WcrBox &bx, const WcrImage &im,
                                              int chunk, int lane)
{
    ChunkCtx c;
    c.K = {im.fx, im.fy, im.cx, im.cy};
    clamp_bounds(bx.clamp0, c.lo0, c.hi0);
    clamp_bounds(bx.clamp1, c.lo1, c.hi1);
    c.npts = __ldcg(a.ws.npts + (size_t)b * a.n_chunks + chunk);
    c.cbase = (size_t)b * a.N + (size_t)chunk * SC_CHUNK;
    c.chunk_n = min(SC_CHUNK, a.N - chunk * SC_CHUNK);
    int area = 0;
    if (a.sc.e_seg > 0 && a.seg_tight) {
        area = __ldg(a.ws.mint + (size_t)b * ((MG + 1) * (MG + 1)) + (MG + 1) * (MG + 1) - 1);   // the mask grid's total
    } else if (a.sc.e_seg > 0) {
        const Roi r1 = box_roi(bx, im.H, im.W, 1, true);
        const int nrows = r1.ncols > 0 ? r1.nrows : 0, wpr = (r1.ncols + 31) >> 5;
        const uint32_t *bits = a.seg_bits + a.seg_off[b];
        for (int w = lane; w < nrows * wpr; w += 32) area += __popc(__ldg(bits + w));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) area += __shfl_xor_sync(0xffffffffu, area, o);
    }
    c.area = area;
    return c;
}

// scores (or bounds, with_seg == false) of the whole chunk -> sc[SC_CHUNK] (+ per-proposal outputs)
__device__ __forceinline__ void chunk_scores(const ScoreArgs &a, int b, const WcrBox &bx, const WcrImage &im,
                                             const ChunkCtx &c, int chunk, bool with_seg, float *sc, int lane)
{
    if (a.out_npts && chunk == 0 && lane == 0) a.out_npts[b] = c.npts;
    if (a.siou_ready) {
        // fast path: the IoU term was computed per thread by score_prep_kernel; a lane only combines
        // it with the point count (and the mask ratio when every hull was filled)
        const float mb = a.sc.map_beta;
#pragma unroll 4
        for (int i = 0; i < SC_CHUNK / 32; ++i) {
            const int loc = i * 32 + lane;
            float score = -INFINITY;
            if (loc < c.chunk_n) {
                const size_t o = c.cbase + loc;
                const float siou = __ldcg(a.ws.siou + o);
                const int count = __ldcg(a.ws.count + o);
                const float spts = c.npts > 0 ? (float)count / (float)c.npts : 0.f;
                float sseg = 1.f;
                if (a.sc.e_seg > 0 && with_seg) {
                    const int2 sc2 = __ldcg(a.ws.segcnt + o);
                    sseg = seg_ratio(sc2.x, sc2.y, c.area, a.sc.seg_mode);
                }
                score = (ipow(shape_term(siou, mb), a.sc.e_iou) * ipow(shape_term(spts, mb), a.sc.e_pts)) *
                        ipow(shape_term(sseg, mb), a.sc.e_seg);
            }
            sc[loc] = score;
        }
        __syncwarp();
        return;
    }
#pragma unroll 1
    for (int i = 0; i < SC_CHUNK / 32; ++i) {
        const int loc = i * 32 + lane;
        float score = -INFINITY;
        if (loc < c.chunk_n) {
            const size_t o = c.cbase + loc;
            float cube[15];
            fetch_cube(a.src, im, bx, b, chunk * SC_CHUNK + loc, cube);
            const PropTerms t = prop_terms(a, cube, o, bx, c.K, c.lo0, c.hi0, c.lo1, c.hi1, c.npts, c.area, with_seg);
            score = t.score;
            if (a.out_scores) a.out_scores[o] = t.score;
            if (a.out_iou) a.out_iou[o] = t.siou;
            if (a.out_seg) a.out_seg[o] = t.sseg;
            if (a.out_count) a.out_count[o] = t.count;
            if (a.out_proj) {
                float *d = a.out_proj + o * 4;
                d[0] = t.pb.x1; d[1] = t.pb.y1; d[2] = t.pb.x2; d[3] = t.pb.y2;
            }
        }
        sc[loc] = score;
    }
    __syncwarp();
}

// factor <= 1 that bounds the mask factor of proposal o (1 without bounds)
__device__ __forceinline__ float seg_bound_factor(const ScoreArgs &a, size_t o, int area)
{
    if (!a.seg_tight) return 1.f;
    const int2 sb = __ldg(a.ws.segb + o);
    // |mask & hull| <= the mask pixels of the grid cells the hull's scan window touches
    const unsigned w = __ldg(a.ws.segwin + o);
    const int gx0 = w & 63, gx1 = (w >> 6) & 63, gy0 = (w >> 12) & 63, gy1 = (w >> 18) & 63;
    const int *mi = a.ws.mint + (o / (size_t)a.N) * (size_t)((MG + 1) * (MG + 1));
    const int in_cells = (__ldg(mi + gy1 * (MG + 1) + gx1) - __ldg(mi + gy0 * (MG + 1) + gx1)) -
                         (__ldg(mi + gy1 * (MG + 1) + gx0) - __ldg(mi + gy0 * (MG + 1) + gx0));
    return ipow(seg_ratio_bound(min(sb.x, in_cells), sb.y, area), a.sc.e_seg);
}

// iou^a * pts^b of proposal q of the chunk, the product chunk_scores forms before the mask factor
// (a.seg_tight: siou_ready and map_beta == 0, so this is chunk_scores' fast path bit for bit)
__device__ __forceinline__ float fill_product(const ScoreArgs &a, const ChunkCtx &c, int q)
{
    const float siou = __ldcg(a.ws.siou + c.cbase + q);
    const int count = __ldcg(a.ws.count + c.cbase + q);
    const float spts = c.npts > 0 ? (float)count / (float)c.npts : 0.f;
    return ipow(siou, a.sc.e_iou) * ipow(spts, a.sc.e_pts);
}

// hull fill of up to 32 proposals of the chunk (q = the lane's proposal or -1) whose bound is in
// `ub`; returns the lane's actual score (-inf without a proposal) and stores it to ws.ub
__device__ __forceinline__ float chunk_fill(const ScoreArgs &a, int b, const WcrBox &bx, const WcrImage &im,
                                            const ChunkCtx &c, int chunk, int q, float ub, float *wbuf, int lane,
                                            bool known = false)
{
    // known: ws.segcnt already holds this lane's proposal (filled by the first pass's epilogue)
    const int qf = known ? -1 : q;
    const unsigned have = __ballot_sync(0xffffffffu, qf >= 0);
    if (have) {
        if (lane == 0) atomicAdd(a.ws.hulls, (unsigned long long)__popc(have));
        seg_group(a.src, a.seg_bits + a.seg_off[b], a.ws.segcnt, b, bx, im, chunk, qf, wbuf, lane);
    }
    __syncwarp();
    float actual = -INFINITY;
    if (q >= 0) {
        const int2 c2 = __ldcg(a.ws.segcnt + c.cbase + q);  // written by this very lane, or earlier (known)
        const float sseg = seg_ratio(c2.x, c2.y, c.area, a.sc.seg_mode);
        actual = ub * ipow(shape_term(sseg, a.sc.map_beta), a.sc.e_seg);   // the product prop_terms forms
        __stcg(a.ws.ub + c.cbase + q, actual);
    }
    return actual;
}

// top-k of sc[] -> records (or chunk candidates); the last chunk of a box merges
__device__ __forceinline__ void chunk_select(const ScoreArgs &a, int b, const WcrBox &bx, const WcrImage &im,
                                             const ChunkCtx &c, int chunk, float *sc, float *wbuf, int lane)
{
    // k rounds of (max score, lowest index); winner r stays with lane r
    int my_win = -1;
    float my_wsc = 0.f;
    for (int r = 0; r < a.k; ++r) {
        float bvv = -INFINITY; int bi = 0x7fffffff;
#pragma unroll 8
        for (int i = 0; i < SC_CHUNK / 32; ++i) {
            const int loc = i * 32 + lane;
            const float vv = sc[loc];
            if (vv > bvv) { bvv = vv; bi = loc; }   // ascending loc keeps the lowest index
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bvv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bvv || (ov == bvv && oi < bi)) { bvv = ov; bi = oi; }
        }
        if (bvv == -INFINITY) bi = -1;
        if (lane == r) { my_win = bi; my_wsc = bvv; }
        if (bi >= 0 && (bi & 31) == lane) sc[bi] = -INFINITY;   // taken
        __syncwarp();
    }
    if (a.seg_lazy) {
        // a winner with a zero score may never have been evaluated (zero bound): (re)fill its
        // hull so that the record carries its s_seg; refilling an evaluated one changes nothing
        const int q = (lane < a.k && my_win >= 0 && my_wsc == 0.f) ? my_win : -1;
        const unsigned need = __ballot_sync(0xffffffffu, q >= 0);
        if (need) {
            if (lane == 0) atomicAdd(a.ws.hulls, (unsigned long long)__popc(need));
            seg_group(a.src, a.seg_bits + a.seg_off[b], a.ws.segcnt, b, bx, im, chunk, q, wbuf, lane);
        }
        __syncwarp();
    }
    float *rec_base = a.n_chunks == 1
                          ? a.records + (size_t)b * a.k * WCR_RECORD_FLOATS
                          : a.ws.cand + ((size_t)b * a.n_chunks + chunk) * a.k * WCR_RECORD_FLOATS;
    if (lane < a.k) {
        float *rec = rec_base + (size_t)lane * WCR_RECORD_FLOATS;
        if (my_win < 0) {
            rec[0] = __int_as_float(-1);
            for (int i = 1; i < WCR_RECORD_FLOATS; ++i) rec[i] = 0.f;
        } else {
            const int j = chunk * SC_CHUNK + my_win;
            const size_t o = (size_t)b * a.N + j;
            float cube[15];
            fetch_cube(a.src, im, bx, b, j, cube);
            const PropTerms t = prop_terms(a, cube, o, bx, c.K, c.lo0, c.hi0, c.lo1, c.hi1, c.npts, c.area);
            rec[0] = __int_as_float(j);
            rec[1] = my_wsc;
#pragma unroll
            for (int i = 0; i < 15; ++i) rec[2 + i] = cube[i];
            rec[17] = t.siou; rec[18] = t.spts; rec[19] = t.sseg;
        }
    }
    if (a.n_chunks > 1) {
        __threadfence();
        __syncwarp();
        int last = 0;
        if (lane == 0) last = (atomicAdd(a.ws.done_box + b, 1) == a.n_chunks - 1);
        last = __shfl_sync(0xffffffffu, last, 0);
        if (last) {
            __threadfence();
            warp_merge(a.ws.cand + (size_t)b * a.n_chunks * a.k * WCR_RECORD_FLOATS,
                       a.n_chunks * a.k, a.k, a.records + (size_t)b * a.k * WCR_RECORD_FLOATS);
        }
    }
}

// phase B of the on-demand path: every batch of the chunk has retired
__device__ __forceinline__ void chunk_finalize_lazy(const ScoreArgs &a, int b, const WcrBox &bx, const WcrImage &im,
                                                    const ChunkCtx &c, int chunk, float *wbuf, int lane)
{
    float *sc = wbuf + TL_SCORE_OFF;
    if (a.pts_lazy && a.seg_tight) {
        // only the proposals that were tested (first pass + survivors, ws.order) have a value
        const int bc = b * a.n_chunks + chunk;
        const uint16_t *list = a.ws.order + (size_t)bc * SC_CHUNK;
        const int nlive = min(a.pa, c.chunk_n) + __ldcg(a.ws.bready + bc) - 1;
#pragma unroll 8
        for (int i = 0; i < SC_CHUNK / 32; ++i) sc[i * 32 + lane] = -INFINITY;
        __syncwarp();
        for (int s = lane; s < nlive; s += 32) {
            const int loc = (int)__ldcg(list + s);
            sc[loc] = __ldcg(a.ws.ub + c.cbase + loc);
        }
    } else {
#pragma unroll 8
        for (int i = 0; i < SC_CHUNK / 32; ++i) {
            const int loc = i * 32 + lane;
            sc[loc] = loc < c.chunk_n ? __ldcg(a.ws.ub + c.cbase + loc) : -INFINITY;
        }
    }
    __syncwarp();
    chunk_select(a, b, bx, im, c, chunk, sc, wbuf, lane);
}

// the warp that retired the chunk's last static item
__device__ __forceinline__ void chunk_epilogue_warp(const ScoreArgs &a, int b, const WcrBox &bx,
                                                    const WcrImage &im, int chunk, float *wbuf, int lane)
{
    const ChunkCtx c = chunk_ctx(a, b, bx, im, chunk, lane);
    float *sc = wbuf + TL_SCORE_OFF;
    if (!a.seg_lazy) {
        chunk_scores(a, b, bx, im, c, chunk, true, sc, lane);
        chunk_select(a, b, bx, im, c, chunk, sc, wbuf, lane);
        return;
    }
    // ---- phase A ---------------------------------------------------------------------------
    if (a.pts_lazy && a.seg_tight) {
        // Point term on demand: only the proposals that were tested -- the first pass and the
        // survivors, the head of ws.order -- can be among the top-k (every other bound is below k
        // actual scores); the rest is never looked at again (-inf in the selection)
        if (a.out_npts && chunk == 0 && lane == 0) a.out_npts[b] = c.npts;
        const int bc = b * a.n_chunks + chunk;
        const uint16_t *list = a.ws.order + (size_t)bc * SC_CHUNK;
        const int nlive = min(a.pa, c.chunk_n) + __ldcg(a.ws.bready + bc) - 1;
#pragma unroll 8
        for (int i = 0; i < SC_CHUNK / 32; ++i) sc[i * 32 + lane] = -INFINITY;
        __syncwarp();
        for (int s = lane; s < nlive; s += 32) {
            const int loc = (int)__ldcg(list + s);
            const float tight = fill_product(a, c, loc) * seg_bound_factor(a, c.cbase + loc, c.area);
            sc[loc] = tight;
            __stcg(a.ws.ub + c.cbase + loc, tight);
        }
    } else {
        chunk_scores(a, b, bx, im, c, chunk, false, sc, lane);
#pragma unroll 8
        for (int i = 0; i < SC_CHUNK / 32; ++i) {
            const int loc = i * 32 + lane;
            if (loc < c.chunk_n) {
                // with mask-term bounds, ws.ub and sc[] hold product x bound of the mask factor; the
                // fills recompute the plain product (fill_product)
                if (a.seg_tight) sc[loc] = sc[loc] * seg_bound_factor(a, c.cbase + loc, c.area);
                __stcg(a.ws.ub + c.cbase + loc, sc[loc]);
            }
        }
    }
    __syncwarp();
    unsigned taken = 0u;   // bit i <-> loc i*32+lane: evaluated before the candidates are cut
    float T = -INFINITY;
    if (a.pts_lazy && a.seg_tight) {
        // Point term on demand: the first pass's epilogue has already filled the hulls of its 32
        // best proposals (ws.evald) -- the chunk's likeliest winners. Their actual scores give T
        // without a single fill on this warp's critical path; everything else that can still
        // reach T goes to the batches.
        float *kv = wbuf;                                  // hull scratch, idle here
        int nk = 0;
#pragma unroll 1
        for (int i = 0; i < SC_CHUNK / 32; ++i) {
            const unsigned evw = __ldcg(a.ws.evald + ((size_t)b * a.n_chunks + chunk) * 32 + i);   // uniform
            if (evw == 0u) continue;
            const bool kn = (evw >> lane) & 1u;
            float act = -INFINITY;
            if (kn) {
                const int q2 = i * 32 + lane;
                const int2 c2 = __ldcg(a.ws.segcnt + c.cbase + q2);
                const float sseg = seg_ratio(c2.x, c2.y, c.area, a.sc.seg_mode);
                act = fill_product(a, c, q2) * ipow(shape_term(sseg, a.sc.map_beta), a.sc.e_seg);
                __stcg(a.ws.ub + c.cbase + q2, act);
                sc[q2] = act;
                taken |= 1u << i;
            }
            const unsigned bm = __ballot_sync(0xffffffffu, kn);
            if (kn) kv[nk + __popc(bm & ((1u << lane) - 1u))] = act;
            nk += __popc(bm);
        }
        __syncwarp();
        const float mine = lane < nk ? kv[lane] : -INFINITY;   // nk <= 32: one batch was filled
        int rank = 0;
        for (int l = 0; l < 32; ++l) {
            const float v = __shfl_sync(0xffffffffu, mine, l);
            rank += (v > mine) || (v == mine && l < lane);
        }
        const unsigned kth = __ballot_sync(0xffffffffu, rank == a.k - 1 && mine > -INFINITY);
        if (kth) T = __shfl_sync(0xffffffffu, mine, __ffs(kth) - 1);
        __syncwarp();
    } else {
    // the 32 best positive bounds (non-destructive arg-max rounds)
    int q = -1;
    float q_ub = 0.f;
    for (int r = 0; r < 32; ++r) {
        float bvv = 0.f; int bi = 0x7fffffff;
#pragma unroll 8
        for (int i = 0; i < SC_CHUNK / 32; ++i) {
            const float vv = sc[i * 32 + lane];
            if (!((taken >> i) & 1u) && vv > bvv) { bvv = vv; bi = i * 32 + lane; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bvv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bvv || (ov == bvv && oi < bi)) { bvv = ov; bi = oi; }
        }
        if (bi == 0x7fffffff) break;                        // uniform: no positive bound left
        if ((bi & 31) == lane) taken |= 1u << (bi >> 5);
        if (lane == r) { q = bi; q_ub = bvv; }
    }
    if (__any_sync(0xffffffffu, q >= 0)) {
        if (a.seg_tight && q >= 0) q_ub = fill_product(a, c, q);
        const float actual = chunk_fill(a, b, bx, im, c, chunk, q, q_ub, wbuf, lane);
        // T = k-th largest actual score of the first batch (rank by counting; ties by lane)
        int rank = 0;
        for (int l = 0; l < 32; ++l) {
            const float v = __shfl_sync(0xffffffffu, actual, l);
            rank += (v > actual) || (v == actual && l < lane);
        }
        const unsigned kth = __ballot_sync(0xffffffffu, rank == a.k - 1 && actual > -INFINITY);
        if (kth) T = __shfl_sync(0xffffffffu, actual, __ffs(kth) - 1);
    }
    }
    // candidates: not in the first batch, positive bound, bound >= T
    uint16_t *cl = a.ws.candlist + ((size_t)b * a.n_chunks + chunk) * SC_CHUNK;
    int n_cand = 0;
#pragma unroll 1
    for (int i = 0; i < SC_CHUNK / 32; ++i) {
        const float vv = sc[i * 32 + lane];
        const bool cnd = !((taken >> i) & 1u) && vv > 0.f && vv >= T;
        const unsigned bm = __ballot_sync(0xffffffffu, cnd);
        if (cnd) cl[n_cand + __popc(bm & ((1u << lane) - 1u))] = (uint16_t)(i * 32 + lane);
        n_cand += __popc(bm);
    }
    // Batches of 32 hulls while tile items remain (a warp walks its 32 hulls one after the other,
    // ~8 us each); once the static queues have run dry the rest of the grid is waiting for exactly
    // these hulls, and batches of 8 put four times as many warps on them.
    const int claimed = *((volatile int *)(a.pts_lazy ? a.ws.queue_b : a.ws.queue));
    const int bs = claimed >= __ldg(a.ws.item_prefix + a.n_boxes) ? 8 : 32;
    const int n_batches = (n_cand + bs - 1) / bs;            // <= 128
    if (n_batches == 0) {
        __threadfence();                                     // first-batch scores were stored by other lanes
        __syncwarp();
        chunk_finalize_lazy(a, b, bx, im, c, chunk, wbuf, lane);
    } else {
        const int bc = b * a.n_chunks + chunk;
        if (lane == 0) { a.ws.ncand[bc] = n_cand | (bs << 16); }
        __threadfence();                                     // bounds, first-batch scores, list, count
        __syncwarp();
        if (lane == 0) {
            const int start = atomicAdd(a.ws.dynq + 1, n_batches);      // reserve
            for (int i = 0; i < n_batches; ++i)
                *((volatile int *)a.ws.segq + start + i) = bc * 128 + i + 1;  // publish (0 = not yet)
        }
    }
    __threadfence();
    __syncwarp();
    if (lane == 0) atomicAdd(a.ws.dynq + 2, 1);              // chunks through phase A
}

// One entry of the dynamic queue: batch `batch` of the candidates of (box, chunk).
__device__ __forceinline__ void dyn_seg_batch(const ScoreArgs &a, int code, float *wbuf, int lane)
{
    const int bc = code >> 7, batch = code & 127;
    const int b = bc / a.n_chunks, chunk = bc - b * a.n_chunks;
    const WcrBox bx = a.boxes[b];
    const WcrImage im = a.images[bx.image];
    const ChunkCtx c = chunk_ctx(a, b, bx, im, chunk, lane);
    const int nc_bs = __ldcg(a.ws.ncand + bc);
    const int n_cand = nc_bs & 0xffff, bs = nc_bs >> 16;     // candidates, hulls per batch (8 or 32)
    const uint16_t *cl = a.ws.candlist + (size_t)bc * SC_CHUNK;
    int q = -1;
    float ub = 0.f;
    if (lane < bs && batch * bs + lane < n_cand) {
        q = (int)__ldcg(cl + batch * bs + lane);
        ub = a.seg_tight ? fill_product(a, c, q) : __ldcg(a.ws.ub + c.cbase + q);
    }
    chunk_fill(a, b, bx, im, c, chunk, q, ub, wbuf, lane);
    __threadfence();
    __syncwarp();
    int last = 0;
    if (lane == 0) last = (atomicAdd(a.ws.seg_pending + bc, 1) == (n_cand + bs - 1) / bs - 1);
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last) return;
    __threadfence();
    chunk_finalize_lazy(a, b, bx, im, c, chunk, wbuf, lane);
}

// ---- point term on demand (a.pts_lazy) -------------------------------------------------
// The tiles of a box are walked TWICE. The first pass tests only the a.pa proposals of each chunk
// with the largest score bound (ws.order, score_order_kernel); the warp that retires the chunk's
// last first-pass item takes T = the k-th best ACTUAL score among them (mask term filled in for
// the 32 best when there is one) and keeps, of the other proposals, those whose bound
//     ub = iou^a * (min(ubcnt, P_b) / P_b)^b          (ubcnt: score_prep_kernel)
// is not below T. Every factor of the score is rounded the way the bound's factors are, the
// count is <= ubcnt and the mask factor <= 1, so score <= ub in floating point: a proposal with
// ub < T has k scores above it and cannot enter the top-k, not even through a tie. It is never
// tested against the depth points (its count stays 0, its stored score 0 < T). The survivors are
// compacted behind the first a.pa entries of ws.order; the second pass (items of a second queue
// that open once every first-pass item has been claimed) tests them, and the warp that retires the
// chunk's last second-pass item runs the ordinary epilogue.
__device__ __forceinline__ float bound_score(const ScoreArgs &a, float siou, int count, int npts)
{
    const float spts = npts > 0 ? (float)count / (float)npts : 0.f;
    return ipow(siou, a.sc.e_iou) * ipow(spts, a.sc.e_pts);     // map_beta == 0: shape_term is the identity
}

__device__ __forceinline__ void chunk_epilogue_first(const ScoreArgs &a, int b, const WcrBox &bx,
                                                     const WcrImage &im, int chunk, float *wbuf, int lane)
{
    const ChunkCtx c = chunk_ctx(a, b, bx, im, chunk, lane);
    const int bc = b * a.n_chunks + chunk;
    uint16_t *list = a.ws.order + (size_t)bc * SC_CHUNK;
    const int PA = min(a.pa, c.chunk_n);
    float *sc = wbuf + TL_SCORE_OFF;                 // [128] first-pass scores with the mask factor's bound
    int *sl = reinterpret_cast<int *>(sc + 128);     // [128] their chunk-relative indices
    float *pu = sc + 256;                            // [128] ... without any mask factor
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int s = i * 32 + lane;
        float v = -INFINITY, pv = -INFINITY; int loc = 0;
        if (s < PA) {
            loc = (int)__ldcg(list + s);
            pv = bound_score(a, __ldcg(a.ws.siou + c.cbase + loc), __ldcg(a.ws.count + c.cbase + loc), c.npts);
            v = a.sc.e_seg > 0 ? pv * seg_bound_factor(a, c.cbase + loc, c.area) : pv;
        }
        sc[s] = v; sl[s] = loc; pu[s] = pv;
    }
    __syncwarp();
    float T = -INFINITY;
    if (a.sc.e_seg > 0) {
        // the 32 best positive values get their hull filled; T = k-th best actual score among them
        unsigned taken = 0u;
        int q = -1;
        float q_ub = 0.f;
        for (int r = 0; r < 32; ++r) {
            float bvv = 0.f; int bi = 0x7fffffff;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float vv = sc[i * 32 + lane];
                if (!((taken >> i) & 1u) && vv > bvv) { bvv = vv; bi = i * 32 + lane; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bvv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bvv || (ov == bvv && oi < bi)) { bvv = ov; bi = oi; }
            }
            if (bi == 0x7fffffff) break;
            if ((bi & 31) == lane) taken |= 1u << (bi >> 5);
            if (lane == r) { q = sl[bi]; q_ub = pu[bi]; }
        }
        if (__any_sync(0xffffffffu, q >= 0)) {
            const float actual = chunk_fill(a, b, bx, im, c, chunk, q, q_ub, wbuf, lane);
            // the final epilogue will not fill these hulls again
            if (q >= 0) atomicOr(a.ws.evald + (size_t)bc * 32 + (q >> 5), 1u << (q & 31));
            int rank = 0;
            for (int l = 0; l < 32; ++l) {
                const float v = __shfl_sync(0xffffffffu, actual, l);
                rank += (v > actual) || (v == actual && l < lane);
            }
            const unsigned kth = __ballot_sync(0xffffffffu, rank == a.k - 1 && actual > -INFINITY);
            if (kth) T = __shfl_sync(0xffffffffu, actual, __ffs(kth) - 1);
        }
    } else {
        // no mask factor: the values ARE the scores; T = the k-th largest of them
        unsigned taken = 0u;
        for (int r = 0; r < a.k; ++r) {
            float bvv = -INFINITY; int bi = 0x7fffffff;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float vv = sc[i * 32 + lane];
                if (!((taken >> i) & 1u) && vv > bvv) { bvv = vv; bi = i * 32 + lane; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bvv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bvv || (ov == bvv && oi < bi)) { bvv = ov; bi = oi; }
            }
            if (bi == 0x7fffffff) { T = -INFINITY; break; }        // fewer than k comparable scores
            if ((bi & 31) == lane) taken |= 1u << (bi >> 5);
            T = bvv;
        }
    }
    // survivors of the rest, compacted in place behind the first pass (a round's writes stay at or
    // below the slots it has read)
    int nB = 0;
#pragma unroll 1
    for (int base = PA; base < c.chunk_n; base += 32) {
        const int s = base + lane;
        bool kp = false;
        int loc = 0;
        if (s < c.chunk_n) {
            loc = (int)__ldcg(list + s);
            const float siou = __ldcg(a.ws.siou + c.cbase + loc);
            const int ubc = min(__ldcg(a.ws.ubcnt + c.cbase + loc), c.npts);
            float ub = bound_score(a, siou, ubc, c.npts);
            if (a.sc.e_seg > 0) ub = ub * seg_bound_factor(a, c.cbase + loc, c.area);
            kp = !(siou >= 0.f && ub < T);                         // NaN anywhere: kept
        }
        const unsigned bm = __ballot_sync(0xffffffffu, kp);
        if (kp) list[PA + nB + __popc(bm & ((1u << lane) - 1u))] = (uint16_t)loc;
        nB += __popc(bm);
    }
    if (lane == 0 && c.chunk_n - PA - nB > 0) atomicAdd(a.ws.pruned, (unsigned long long)(c.chunk_n - PA - nB));
    __threadfence();                                               // the list, before the flag
    __syncwarp();
    if (lane == 0) *((volatile int *)a.ws.bready + bc) = nB + 1;
}

// Phase attribution for tools/diag_score.py (a separate build with -DWCR_PROFILE_PHASES, never the
// product): SM cycles per phase, summed over warps, in the workspace header at byte 64 + 8 * i.
#ifdef WCR_PROFILE_PHASES
#define PH_T0() const long long ph_t0_ = clock64()
#define PH_MARK(v) v = clock64()
#define PH_ADDV(i, v) do { if (lane == 0) atomicAdd((unsigned long long *)((char *)a.ws.queue + 64) + (i), \
                                                    (unsigned long long)(clock64() - (v))); } while (0)
#define PH_ADD(i) do { if (lane == 0) atomicAdd((unsigned long long *)((char *)a.ws.queue + 64) + (i), \
                                                (unsigned long long)(clock64() - ph_t0_)); } while (0)
#else
#define PH_T0() do { } while (0)
#define PH_ADD(i) do { } while (0)
#define PH_MARK(v) do { } while (0)
#define PH_ADDV(i, v) do { } while (0)
#endif

template <int WARPS, int MINB, int GRP, bool F2, bool ICNT = false>
__global__ void __launch_bounds__(WARPS * 32, MINB) score_tiles_kernel(const ScoreArgs a)
{
    __shared__ __align__(16) float s_wbuf[WARPS][TL_WBUF];   // per-warp scratch: hulls / scores
    __shared__ int s_prefix[2048];                            // item prefix of up to 2047 boxes: the decode's
                                                              // binary search then never leaves the SM
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    float *wbuf = s_wbuf[wid];
    const int *prefix = a.ws.item_prefix;
    if (a.n_boxes + 1 <= 2048) {
        for (int i = threadIdx.x; i <= a.n_boxes; i += WARPS * 32) s_prefix[i] = a.ws.item_prefix[i];
        __syncthreads();                                      // the only CTA-wide barrier, before any warp diverges
        prefix = s_prefix;
    }
    const bool use_pts = a.sc.e_pts > 0, use_seg = a.sc.e_seg > 0;
    const int total = prefix[a.n_boxes];
    const int subs = SC_CHUNK / a.ps;   // tile items per (chunk, tile)

    // Queues: tile items and seg items have their own counters; the last warp of every CTA prefers
    // seg items, the others tile items, and each falls back to the other queue when its own runs
    // dry -- the two kinds of item run concurrently on an SM (integer-heavy next to FMA-heavy) but
    // a warp rarely switches code path, which the instruction caches reward. Seg items exist only
    // when every mask term is wanted (a.seg_lazy == 0); otherwise the epilogues fill the few hulls
    // they need and every warp works on tiles.
    const int nseg = (use_seg && !a.seg_lazy) ? TL_SEGS : 0;
    const int seg_total = a.n_boxes * a.n_chunks * nseg;
    const bool pref_seg = nseg > 0 && wid == WARPS - 1;
    bool seg_left = seg_total > 0, tile_left = true, pass2_left = a.pts_lazy != 0;
    const int total_chunks = a.n_boxes * a.n_chunks;
    volatile int *dynq = a.ws.dynq;   // [0] tickets drawn, [1] tail (reserved), [2] chunks through phase A
    const long long seg_cap_ll = (long long)total_chunks * 128;
    const int seg_cap = seg_cap_ll > 0x7fffffffLL ? 0x7fffffff : (int)seg_cap_ll;   // entries of ws.segq
    int my_ticket = -1;               // lane 0: queue index this warp owns (-1: none)
    long long ph_loop = 0;
    (void)ph_loop;
#ifdef WCR_PROFILE_PHASES
    unsigned long long ph_start;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ph_start));
    if (lane == 0) atomicMin((unsigned long long *)((char *)a.ws.queue + 64) + 12, ph_start);
#endif
    for (;;) {
        PH_MARK(ph_loop);
        if (a.seg_lazy) {
            // Published hull batches first: they unblock a chunk's top-k. Claims are TICKETS: a warp
            // without one draws the next queue index (one atomicAdd, never retried) and keeps it
            // until that entry is published -- possibly later, the warp does static work meanwhile
            // and looks again after every item. Every index below the final tail thus has exactly
            // one owner, nobody spins on a shared word (a compare-and-swap on the head made
            // every loser retry: 0.44 ms per warp on C2, profiles/variants_tiles_r2.txt).
            int code = 0;
            if (lane == 0) {
                if (my_ticket < 0) my_ticket = atomicAdd(a.ws.dynq, 1);
                if (my_ticket < seg_cap) code = *((volatile int *)a.ws.segq + my_ticket);
                if (code > 0) my_ticket = -1;
            }
            code = __shfl_sync(0xffffffffu, code, 0);
            if (code > 0) {
                __threadfence();
                PH_T0();
                dyn_seg_batch(a, code - 1, wbuf, lane);
                __syncwarp();
                PH_ADD(4);
                continue;
            }
        }
        int item = -1, kind = -1;   // kind: 0 tile queue, 1 seg queue, 2 second pass of the tiles
        while (kind < 0 && (seg_left || tile_left)) {
            const bool try_seg = seg_left && (pref_seg || !tile_left);
            if (lane == 0) item = atomicAdd(a.ws.queue + (try_seg ? 1 : 0), 1);
            item = __shfl_sync(0xffffffffu, item, 0);
            if (try_seg) { if (item < seg_total) kind = 1; else seg_left = false; }
            else { if (item < total) kind = 0; else tile_left = false; }
        }
        if (kind < 0 && pass2_left) {
            // every first-pass item has been claimed (by warps that never wait): the second pass
            // may wait for a chunk's first-pass epilogue without risking a deadlock
            if (lane == 0) item = atomicAdd(a.ws.queue_b, 1);
            item = __shfl_sync(0xffffffffu, item, 0);
            if (item < total) kind = 2; else pass2_left = false;
        }
        if (kind < 0) {
            if (!a.seg_lazy) break;
            // no static item left: serve the dynamic queue until every chunk is through phase A
            // (the tail is final then) and the queue is drained
            int done = 0;
            if (lane == 0) {
                const int pa = dynq[2];
                __threadfence();
                // the tail is final once every chunk is through phase A (a chunk reserves its
                // entries before it reports): a ticket at or beyond it will never be published
                done = pa == total_chunks && my_ticket >= dynq[1];
                if (!done) __nanosleep(100);
            }
            const int dn = __shfl_sync(0xffffffffu, done, 0);
            PH_ADDV(8, ph_loop);   // tail: nothing to claim, batches may still be published
            if (dn) break;
            continue;
        }
        int b = 0, pos = 0;   // box, and its position in the serving order (tile queues)
        if (kind == 1) {
            b = item / (a.n_chunks * TL_SEGS);
        } else {
            int hi = a.n_boxes;   // prefix[pos] <= item < prefix[hi]
            while (hi - pos > 1) {
                const int mid = (pos + hi) >> 1;
                if (prefix[mid] <= item) pos = mid; else hi = mid;
            }
            b = __ldg(a.ws.box_order + pos);
        }
        const WcrBox bx = a.boxes[b];
        const WcrImage im = a.images[bx.image];
        const TileGeom tg = tile_geom(bx, im.H, im.W, a.sc.point_stride, use_pts, GRP);
        const int pt = tile_items_per_chunk(tg, false, subs);
        const int per = pt + nseg;   // items that retire before the chunk's epilogue
        int chunk, seg_grp = -1, ti = 0;
        if (kind == 1) {
            const int r = item - b * (a.n_chunks * TL_SEGS);
            chunk = r / TL_SEGS; seg_grp = r - chunk * TL_SEGS;
        } else {
            const int local = item - prefix[pos];
            chunk = local / pt; ti = local - chunk * pt;
        }
        const int bc = b * a.n_chunks + chunk;
        PH_ADDV(6, ph_loop);       // claim + decode
        const uint16_t *list = nullptr;
        int n_list = 0;
        if (a.pts_lazy) {
            list = a.ws.order + (size_t)bc * SC_CHUNK;
            n_list = min(a.pa, min(SC_CHUNK, a.N - chunk * SC_CHUNK));
            if (kind == 2) {
                int v = 0;
                PH_T0();
                if (lane == 0) {
                    volatile int *f = (volatile int *)a.ws.bready + bc;
                    while ((v = *f) == 0) __nanosleep(64);
                    __threadfence();
                }
                v = __shfl_sync(0xffffffffu, v, 0);
                PH_ADD(5);
                list += n_list; n_list = v - 1;
            }
        }
        if (seg_grp >= 0) {
            seg_item(a, b, bx, im, chunk, seg_grp, wbuf, lane);
        } else {
            const int tile = ti / subs, sub = ti - tile * subs;
            PH_T0();
            if (tile < tg.n_tiles)
                tile_item<GRP, F2, ICNT>(a, b, im, tg, chunk, tile, sub, wbuf, lane, list, n_list, kind != 2);
            PH_ADD(kind == 2 ? 1 : 0);
        }
        // ---- retire the item ------------------------------------------------------
        PH_MARK(ph_loop);
        __threadfence();
        __syncwarp();
        int last = 0;
        int *done = (kind == 2 ? a.ws.done_b : a.ws.done_chunk) + bc;
        if (lane == 0) last = (atomicAdd(done, 1) == per - 1);
        last = __shfl_sync(0xffffffffu, last, 0);
        PH_ADDV(7, ph_loop);
        if (!last) continue;
        __threadfence();
        PH_T0();
        if (a.pts_lazy && kind == 0) chunk_epilogue_first(a, b, bx, im, chunk, wbuf, lane);
        else chunk_epilogue_warp(a, b, bx, im, chunk, wbuf, lane);
        __syncwarp();
        PH_ADD((a.pts_lazy && kind == 0) ? 2 : 3);
    }
#ifdef WCR_PROFILE_PHASES
    unsigned long long ph_end;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ph_end));
    if (lane == 0) {   // [10] latest exit, [11] sum of exits (ns, absolute), [12] earliest start
        atomicMax((unsigned long long *)((char *)a.ws.queue + 64) + 10, ph_end);
        atomicAdd((unsigned long long *)((char *)a.ws.queue + 64) + 11, ph_end - ph_start);
    }
#endif
}

}  // namespace wcr
