/* TEST INFRASTRUCTURE ONLY (oracle). Never linked or loaded by the product path.
 *
 * Plain-C restatement of the pieces of the hot path whose bit pattern depends
 * on where a fused multiply-add sits. Compiled with -ffp-contract=off so the
 * only fused operations are the explicit fmaf() calls below.
 *
 * Each function cites the reference lines it follows (paths relative to
 * /root/reference). Where the reference is torch, the rounding order stated
 * here was established by running the reference in the authoring container
 * (torch 2.11 CPU) -- see tests/golden/make_golden.py and DESIGN.md "Oracle".
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>

/* ProposalNetwork/utils/plane.py:96-113 -- hypothesis plane through 3 sampled
 * points. torch.cross(A,B) on CPU evaluates a1*b2 - a2*b1 as
 * fma(a1, b2, -(a2*b1)); torch.norm(p=2) over 3 values is
 * sqrt(fma(z,z,fma(y,y,x*x))); k = -((c0*q0 + c1*q1) + c2*q2) with q = sample
 * point 1 (plane.py:112), products rounded separately. */
static inline float cross_c(float a1, float b2, float a2, float b1)
{
    float t = a2 * b1;
    return fmaf(a1, b2, -t);
}

void orc_plane_hypotheses(const float *pts, int64_t n, const int64_t *ids,
                          int64_t iters, float *eqs)
{
    (void)n;
    for (int64_t i = 0; i < iters; ++i) {
        const float *p0 = pts + 3 * ids[3 * i + 0];
        const float *p1 = pts + 3 * ids[3 * i + 1];
        const float *p2 = pts + 3 * ids[3 * i + 2];
        float ax = p1[0] - p0[0], ay = p1[1] - p0[1], az = p1[2] - p0[2];
        float bx = p2[0] - p0[0], by = p2[1] - p0[1], bz = p2[2] - p0[2];
        float cx = cross_c(ay, bz, az, by);
        float cy = cross_c(az, bx, ax, bz);
        float cz = cross_c(ax, by, ay, bx);
        float nrm = sqrtf(fmaf(cz, cz, fmaf(cy, cy, cx * cx)));
        cx = cx / nrm; cy = cy / nrm; cz = cz / nrm;
        float k = -((cx * p1[0] + cy * p1[1]) + cz * p1[2]);
        eqs[4 * i + 0] = cx; eqs[4 * i + 1] = cy; eqs[4 * i + 2] = cz; eqs[4 * i + 3] = k;
    }
}

/* plane.py:117-126 -- dist = (((a*x + b*y) + c*z) + d) / sqrt((a*a + b*b) + c*c),
 * inlier iff |dist| <= thresh. Elementwise torch kernels: every op rounded. */
void orc_plane_counts(const float *pts, int64_t n, const float *eqs, int64_t iters,
                      float thresh, int64_t *counts)
{
    for (int64_t i = 0; i < iters; ++i) {
        float a = eqs[4 * i], b = eqs[4 * i + 1], c = eqs[4 * i + 2], d = eqs[4 * i + 3];
        float den = sqrtf((a * a + b * b) + c * c);
        int64_t cnt = 0;
        for (int64_t j = 0; j < n; ++j) {
            float x = pts[3 * j], y = pts[3 * j + 1], z = pts[3 * j + 2];
            float dist = (((a * x + b * y) + c * z) + d) / den;
            cnt += (fabsf(dist) <= thresh);
        }
        counts[i] = cnt;
    }
}
