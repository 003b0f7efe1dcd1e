// iou3d.cu -- exact intersection volume / IoU of oriented 3D boxes.
//
// Replaces pytorch3d.ops.box3d_overlap as the reference calls it: iou_3d
// (ProposalNetwork/utils/utils.py:189-205), roi_heads.py:942, omni3d_evaluation.py:106-166.
// pytorch3d is not vendored in the reference (requirements.txt pins no version); its
// published algorithm is restated: the faces of each box are clipped against the half-spaces
// of the other (Sutherland-Hodgman), the surviving polygons bound the intersection
// polyhedron and its volume follows from the divergence theorem. Corner order and face
// lists are pytorch3d's (_box_planes). Differences kept deliberately small:
//   * faces are clipped as quads, not as two triangles each (same polygons, fewer of them);
//   * coplanar faces with the same outward normal are counted once by a symmetric
//     predicate instead of pytorch3d's epsilon de-duplication.
// One (box1, box2) pair per 16 lanes: lane f < 6 clips face f of box1 against box2, lane
// 6 <= f < 12 clips face f - 6 of box2 against box1; a 16-lane shuffle reduction sums the
// cone volumes. FP32, ALU bound, no roofline claim (a few thousand FLOP per pair).
#include "common.cuh"

namespace wcr {

__constant__ int c_box_face[6][4] = {{0, 1, 2, 3}, {3, 2, 6, 7}, {0, 1, 5, 4},
                                     {0, 3, 7, 4}, {1, 2, 6, 5}, {4, 5, 6, 7}};

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r = {x, y, z}; return r; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b)
{
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ V3 ld3(const float *c, int i) { return v3(c[3 * i], c[3 * i + 1], c[3 * i + 2]); }

struct Plane { V3 n; float d; };   // inside: n . x <= d, |n| = 1 (or 0 for a degenerate face)

__device__ __forceinline__ V3 box_centre(const float *c)
{
    V3 s = v3(0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) s = s + ld3(c, i);
    return s * 0.125f;
}

// plane of face f of a box with outward normal (away from the box centre)
__device__ __forceinline__ Plane face_plane(const float *c, V3 ctr, int f)
{
    const V3 v0 = ld3(c, c_box_face[f][0]), v1 = ld3(c, c_box_face[f][1]), v3_ = ld3(c, c_box_face[f][3]);
    V3 n = cross(v1 - v0, v3_ - v0);
    const float len = sqrtf(dot(n, n));
    n = len > 0.f ? n * (1.f / len) : v3(0.f, 0.f, 0.f);
    if (dot(n, ctr - v0) > 0.f) n = n * -1.f;
    Plane p;
    p.n = n; p.d = dot(n, v0);
    return p;
}

// area of a planar polygon (fan from vertex 0)
__device__ __forceinline__ float poly_area(const V3 *p, int m)
{
    V3 s = v3(0.f, 0.f, 0.f);
    for (int i = 1; i + 1 < m; ++i) s = s + cross(p[i] - p[0], p[i + 1] - p[0]);
    return 0.5f * sqrtf(dot(s, s));
}

// volume of a box from its own six faces (divergence theorem about its centre)
__device__ __forceinline__ float box_volume(const float *c, V3 ctr)
{
    float v = 0.f;
#pragma unroll
    for (int f = 0; f < 6; ++f) {
        V3 q[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) q[k] = ld3(c, c_box_face[f][k]);
        const Plane p = face_plane(c, ctr, f);
        v += poly_area(q, 4) * (p.d - dot(p.n, ctr));
    }
    return v * (1.f / 3.f);
}

// face fx of box X and plane py of box Y lie in one plane with the same outward normal.
// Evaluated by both lanes concerned from the same operands in the same order, so they agree.
__device__ __forceinline__ bool coplanar_same(const float *cx, int fx, const Plane &px, const Plane &py, float eps)
{
    if (!(dot(px.n, py.n) > 0.999f)) return false;
    float worst = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) worst = fmaxf(worst, fabsf(dot(py.n, ld3(cx, c_box_face[fx][k])) - py.d));
    return worst <= eps;
}

__global__ void __launch_bounds__(256)
box3d_overlap_kernel(const float *__restrict__ boxes1, int N, const float *__restrict__ boxes2, int M,
                     float *__restrict__ out_vol, float *__restrict__ out_iou)
{
    const int lane = threadIdx.x & 15;
    const long long pair = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
    const bool live = pair < (long long)N * M;
    const int i = live ? (int)(pair / M) : 0, j = live ? (int)(pair - (long long)i * M) : 0;
    const float *c1 = boxes1 + 24 * (size_t)i, *c2 = boxes2 + 24 * (size_t)j;
    const V3 ctr1 = box_centre(c1), ctr2 = box_centre(c2);
    float contrib = 0.f;
    if (live && lane < 12) {
        const bool first = lane < 6;
        const int f = first ? lane : lane - 6;
        const float *cx = first ? c1 : c2, *cy = first ? c2 : c1;    // X owns the face, Y clips it
        const V3 cX = first ? ctr1 : ctr2, cY = first ? ctr2 : ctr1;
        const Plane own = face_plane(cx, cX, f);
        // length scale for the coplanarity test: the larger box diagonal
        const float scale = sqrtf(fmaxf(dot(ld3(c1, 6) - ld3(c1, 0), ld3(c1, 6) - ld3(c1, 0)),
                                        dot(ld3(c2, 6) - ld3(c2, 0), ld3(c2, 6) - ld3(c2, 0))));
        const float eps = 1e-5f * scale;
        V3 a[12], b[12];
        int m = 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) a[k] = ld3(cx, c_box_face[f][k]);
        bool drop = false;
        for (int g = 0; g < 6 && m > 0; ++g) {
            const Plane p = face_plane(cy, cY, g);
            // coplanar, same normal: box1's face keeps the polygon, box2's copy is dropped.
            // The predicate is always evaluated as (box1 face, box2 plane).
            const bool same = first ? coplanar_same(c1, f, own, p, eps)
                                    : coplanar_same(c1, g, p, own, eps);
            if (same) {
                if (!first) { drop = true; break; }
                continue;                                             // do not clip by the twin plane
            }
            int mo = 0;
            for (int k = 0; k < m; ++k) {
                const V3 s = a[k], e = a[k + 1 == m ? 0 : k + 1];
                const float ds = dot(p.n, s) - p.d, de = dot(p.n, e) - p.d;
                if (ds <= 0.f) {
                    b[mo++] = s;
                    if (de > 0.f) b[mo++] = s + (e - s) * (ds / (ds - de));
                } else if (de <= 0.f) {
                    b[mo++] = s + (e - s) * (ds / (ds - de));
                }
            }
            m = mo;
            for (int k = 0; k < m; ++k) a[k] = b[k];
        }
        if (!drop && m >= 3) contrib = poly_area(a, m) * (own.d - dot(own.n, ctr1)) * (1.f / 3.f);
    }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
    if (live && lane == 0) {
        const float v1 = box_volume(c1, ctr1), v2 = box_volume(c2, ctr2);
        const float inter = fminf(fmaxf(contrib, 0.f), fminf(v1, v2));
        out_vol[pair] = inter;
        out_iou[pair] = inter / ((v1 + v2) - inter);
    }
}

// volume of each box from its own faces (pytorch3d rejects boxes whose volume is ~0)
__global__ void __launch_bounds__(128)
box3d_volume_kernel(const float *__restrict__ boxes, int n, float *__restrict__ vol)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *c = boxes + 24 * (size_t)i;
    vol[i] = box_volume(c, box_centre(c));
}

}  // namespace wcr

extern "C" int wcr_box3d_overlap(const float *boxes1, int N, const float *boxes2, int M, float *out_vol,
                                 float *out_iou, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(boxes1 && boxes2 && out_vol && out_iou, "wcr_box3d_overlap: null pointer");
    WCR_REQUIRE(N >= 0 && M >= 0, "wcr_box3d_overlap: bad sizes");
    const long long pairs = (long long)N * M;
    if (pairs == 0) return WCR_OK;
    const long long blocks = (pairs * 16 + 255) / 256;
    WCR_REQUIRE(blocks <= 2147483647LL, "wcr_box3d_overlap: too many pairs");
    wcr::box3d_overlap_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(boxes1, N, boxes2, M, out_vol,
                                                                               out_iou);
    WCR_CHECK_LAUNCH("box3d_overlap_kernel");
    return WCR_OK;
}

extern "C" int wcr_box3d_volume(const float *boxes, int n, float *out_vol, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(boxes && out_vol, "wcr_box3d_volume: null pointer");
    WCR_REQUIRE(n >= 0, "wcr_box3d_volume: bad size");
    if (n == 0) return WCR_OK;
    wcr::box3d_volume_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(boxes, n, out_vol);
    WCR_CHECK_LAUNCH("box3d_volume_kernel");
    return WCR_OK;
}
