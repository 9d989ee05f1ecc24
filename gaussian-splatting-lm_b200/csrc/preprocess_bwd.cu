// =============================================================================
// preprocess_bwd.cu -- per-Gaussian adjoint stage (a14), one thread per Gaussian.
//
// Replaces computeCov2DCUDA (backward.cu:147-326) + bwd preprocessCUDA (backward.cu:398-449)
// + bwd computeColorFromSH (:23-142) + bwd computeCov3D (:330-393), and the ten float
// atomicAdd per (pixel, Gaussian) hit of the reference's bwd renderCUDA (:593-635):
// the tile kernels leave ONE 48-byte partial per (tile, Gaussian) instance at the instance's
// slot of the per-Gaussian CSR (no atomics); this kernel sums a Gaussian's compact instances -- one contiguous
// run [g_off[i], g_off[i+1]) -- in a fixed order, so the whole backward pass is deterministic.
// Quirks of the reference kept on purpose: 1/(det^2 + 1e-7) in the
// conic adjoint, no quaternion-normalisation Jacobian, the alpha-clamp is not masked.
// =============================================================================
#include "gslm_common.cuh"
#include "gslm_math.cuh"
#include <cstring>

namespace gslm {

struct BwdArgs {
    int P, D, M, W, H;
    float scale_modifier, tan_fovx, tan_fovy;
    int antialiasing, has_invdepth;
    const float *means3D, *opacities, *scales, *rotations, *cov3D, *shs;
    const float *viewmatrix, *projmatrix, *campos;
    const int32_t* radii;
    const uint8_t* clamped;
    const uint32_t* g_off;
    const uint32_t* active_ids;   // accumulate mode: only the Gaussians with compact instances are visited
    const uint32_t* n_active;
    const float4* inst_grads;
    GslmGrads g;
};

template <bool ACC> __device__ __forceinline__ void put(float* p, size_t i, float v) {
    if (p) {
        if (ACC) p[i] += v; else p[i] = v;
    }
}

// the whole per-Gaussian adjoint of Gaussian idx for the view described by (a, sV, sPV, sCam)
// PHASED (the dense multi-view kernel): lanes of one warp may hold the SAME Gaussian for different views.  Everything
// up to the two store sites runs concurrently; at each store site the lanes take turns in view order (`rank` among the
// lanes with equal idx, `maxrank` its warp maximum, `callmask` the lanes inside this call), so the read-modify-write
// sequence on a row is exactly the view-by-view sequence of the unphased kernel.
template <bool ACC, bool PHASED = false>
__device__ __forceinline__ void bwd_gaussian(const BwdArgs& a, const float* __restrict__ sV, const float* __restrict__ sPV,
                                             const float* __restrict__ sCam, int idx, uint32_t callmask = 0xffffffffu,
                                             int rank = 0, int maxrank = 0) {
    const size_t i = idx;
    const int ncoef = a.shs ? a.M : 0;
    const uint32_t beg = a.g_off[idx], end = a.g_off[idx + 1];
    uint32_t syncmask = callmask;
    if (PHASED) syncmask = __ballot_sync(callmask, !(a.radii[idx] <= 0 || beg == end));
    // culled, or visible but blended by no pixel: every gradient is exactly zero
    if (a.radii[idx] <= 0 || beg == end) {
        if (!ACC) {  // overwrite mode: culled Gaussians get exact zeros
            for (int k = 0; k < 3; k++) { put<false>(a.g.dL_dmeans2D, 3 * i + k, 0.f); put<false>(a.g.dL_dcolors, 3 * i + k, 0.f);
                                          put<false>(a.g.dL_dmeans3D, 3 * i + k, 0.f); put<false>(a.g.dL_dscales, 3 * i + k, 0.f); }
            put<false>(a.g.dL_dopacity, i, 0.f);
            for (int k = 0; k < 6; k++) put<false>(a.g.dL_dcov3D, 6 * i + k, 0.f);
            for (int k = 0; k < 4; k++) put<false>(a.g.dL_drotations, 4 * i + k, 0.f);
            if (a.g.dL_dsh && ncoef) {
                float* row = a.g.dL_dsh + (size_t)3 * ncoef * i;
                if ((ncoef & 3) == 0 && (reinterpret_cast<size_t>(a.g.dL_dsh) & 15u) == 0) {   // 16-byte aligned rows
                    float4* r4 = reinterpret_cast<float4*>(row);
                    for (int c = 0; c < 3 * ncoef / 4; c++) r4[c] = make_float4(0.f, 0.f, 0.f, 0.f);
                } else {
                    for (int k = 0; k < 3 * ncoef; k++) row[k] = 0.f;
                }
            }
        }
        return;
    }
    // ---- 1. deterministic sum of this Gaussian's instance partials: the tile kernels wrote them at the CSR slots
    //         [beg, end), i.e. one contiguous run in ascending tile order ----
    float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
    float2 s2 = make_float2(0.f, 0.f);
    {
        const float4* r = a.inst_grads + (size_t)REC_F4 * beg;
        uint32_t slot = beg;
        // four records in flight per round trip (the adds keep their order: same bits); the run of a Gaussian is
        // ~13 records long and one dependent load per record made this loop the latency of the whole stage
        for (; slot + 4 <= end; slot += 4, r += 4 * REC_F4) {
            float4 q[4][3];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                q[j][0] = __ldg(r + REC_F4 * j); q[j][1] = __ldg(r + REC_F4 * j + 1); q[j][2] = __ldg(r + REC_F4 * j + 2);
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                s0.x += q[j][0].x; s0.y += q[j][0].y; s0.z += q[j][0].z; s0.w += q[j][0].w;
                s1.x += q[j][1].x; s1.y += q[j][1].y; s1.z += q[j][1].z; s1.w += q[j][1].w;
                s2.x += q[j][2].x; s2.y += q[j][2].y;
            }
        }
        for (; slot < end; slot++, r += REC_F4) {
            const float4 q0 = __ldg(r), q1 = __ldg(r + 1), q2 = __ldg(r + 2);
            s0.x += q0.x; s0.y += q0.y; s0.z += q0.z; s0.w += q0.w;
            s1.x += q1.x; s1.y += q1.y; s1.z += q1.z; s1.w += q1.w;
            s2.x += q2.x; s2.y += q2.y;
        }
    }
    const float g2x = s0.x, g2y = s0.y;                 // dL/dmean2D (NDC-scaled, backward.cu:527-528)
    const float dcx = s0.z, dcy = s0.w, dcz = s1.x;     // dL/dconic (xx, xy, yy)
    float dL_dop = s1.y;
    float dcol[3] = {s1.z, s1.w, s2.x};
    const float dinvd = s2.y;

    // ---- 2. conic -> cov2D -> (cov3D, mean)    backward.cu:147-326 ----
    const float* c3 = a.cov3D + 6 * i;
    const float mx = a.means3D[3 * i], my = a.means3D[3 * i + 1], mz = a.means3D[3 * i + 2];
    float tx = sV[0] * mx + sV[4] * my + sV[8] * mz + sV[12];
    float ty = sV[1] * mx + sV[5] * my + sV[9] * mz + sV[13];
    const float tz = sV[2] * mx + sV[6] * my + sV[10] * mz + sV[14];
    const float fx = a.W / (2.0f * a.tan_fovx), fy = a.H / (2.0f * a.tan_fovy);
    const float limx = 1.3f * a.tan_fovx, limy = 1.3f * a.tan_fovy;
    const float txtz = tx / tz, tytz = ty / tz;
    tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
    ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
    const float xmul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float ymul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    const float J00 = fx / tz, J02 = -(fx * tx) / (tz * tz), J11 = fy / tz, J12 = -(fy * ty) / (tz * tz);
    // rows of the 2x3 projection U = J * Wr (u0 = glm T[0], u1 = glm T[1])
    float u0[3], u1[3];
    const float w0[3] = {sV[0], sV[4], sV[8]}, w1[3] = {sV[1], sV[5], sV[9]}, w2[3] = {sV[2], sV[6], sV[10]};
#pragma unroll
    for (int r = 0; r < 3; r++) { u0[r] = w0[r] * J00 + w2[r] * J02; u1[r] = w1[r] * J11 + w2[r] * J12; }
    const float S[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
    float Su0[3], Su1[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        Su0[r] = u0[0] * S[r][0] + u0[1] * S[r][1] + u0[2] * S[r][2];
        Su1[r] = u1[0] * S[r][0] + u1[1] * S[r][1] + u1[2] * S[r][2];
    }
    float c_xx = u0[0] * Su0[0] + u0[1] * Su0[1] + u0[2] * Su0[2];
    float c_xy = u0[0] * Su1[0] + u0[1] * Su1[1] + u0[2] * Su1[2];
    float c_yy = u1[0] * Su1[0] + u1[1] * Su1[1] + u1[2] * Su1[2];
    const float h_var = 0.3f;
    float d_inside_root = 0.f;
    if (a.antialiasing) {  // backward.cu:208-246
        const float det_cov = c_xx * c_yy - c_xy * c_xy;
        c_xx += h_var; c_yy += h_var;
        const float det_plus = c_xx * c_yy - c_xy * c_xy;
        const float ratio = det_cov / det_plus;
        const float hcs = sqrtf(fmaxf(0.000025f, ratio));
        const float d_hcs = dL_dop * a.opacities[i];
        dL_dop = dL_dop * hcs;
        d_inside_root = ratio <= 0.000025f ? 0.f : d_hcs / (2.f * hcs);
    } else {
        c_xx += h_var; c_yy += h_var;
    }
    float da = 0.f, db = 0.f, dc = 0.f;   // dL/d(c_xx, c_xy, c_yy)
    if (a.antialiasing) {
        const float x = c_xx, y = c_yy, z = c_xy, w = h_var;
        const float dn = w * w + w * (x + y) + x * y - z * z;
        const float f = d_inside_root / (dn * dn);
        da = w * (w * y + y * y + z * z) * f;
        dc = w * (w * x + x * x + z * z) * f;
        db = -2.f * w * z * (w + x + y) * f;
    }
    const float denom = c_xx * c_yy - c_xy * c_xy;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (denom2inv != 0.f) {
        da += denom2inv * (-c_yy * c_yy * dcx + 2.f * c_xy * c_yy * dcy + (denom - c_xx * c_yy) * dcz);
        dc += denom2inv * (-c_xx * c_xx * dcz + 2.f * c_xx * c_xy * dcy + (denom - c_xx * c_yy) * dcx);
        db += denom2inv * 2.f * (c_xy * c_yy * dcx - (denom + 2.f * c_xy * c_xy) * dcy + c_xx * c_xy * dcz);
        dcov[0] = u0[0] * u0[0] * da + u0[0] * u1[0] * db + u1[0] * u1[0] * dc;
        dcov[3] = u0[1] * u0[1] * da + u0[1] * u1[1] * db + u1[1] * u1[1] * dc;
        dcov[5] = u0[2] * u0[2] * da + u0[2] * u1[2] * db + u1[2] * u1[2] * dc;
        dcov[1] = 2.f * u0[0] * u0[1] * da + (u0[0] * u1[1] + u0[1] * u1[0]) * db + 2.f * u1[0] * u1[1] * dc;
        dcov[2] = 2.f * u0[0] * u0[2] * da + (u0[0] * u1[2] + u0[2] * u1[0]) * db + 2.f * u1[0] * u1[2] * dc;
        dcov[4] = 2.f * u0[2] * u0[1] * da + (u0[1] * u1[2] + u0[2] * u1[1]) * db + 2.f * u1[1] * u1[2] * dc;
    }
    // dL/du0 = 2 S u0 da + S u1 db ; dL/du1 = 2 S u1 dc + S u0 db   (backward.cu:277-283)
    float dJ00 = 0.f, dJ02 = 0.f, dJ11 = 0.f, dJ12 = 0.f;
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const float du0 = 2.f * Su0[r] * da + Su1[r] * db;
        const float du1 = 2.f * Su1[r] * dc + Su0[r] * db;
        dJ00 += w0[r] * du0; dJ02 += w2[r] * du0;
        dJ11 += w1[r] * du1; dJ12 += w2[r] * du1;
    }
    const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
    const float dtx = xmul * -fx * itz2 * dJ02;
    const float dty = ymul * -fy * itz2 * dJ12;
    float dtz = -fx * itz2 * dJ00 - fy * itz2 * dJ11 + (2.f * fx * tx) * itz3 * dJ02 + (2.f * fy * ty) * itz3 * dJ12;
    if (a.has_invdepth) dtz -= dinvd / (tz * tz);   // backward.cu:314-315
    float dmean[3] = {sV[0] * dtx + sV[1] * dty + sV[2] * dtz, sV[4] * dtx + sV[5] * dty + sV[6] * dtz,
                      sV[8] * dtx + sV[9] * dty + sV[10] * dtz};

    // ---- 3. projective divide path of mean2D     backward.cu:426-440 ----
    {
        const float hw = sPV[3] * mx + sPV[7] * my + sPV[11] * mz + sPV[15];
        const float m_w = 1.0f / (hw + 0.0000001f);
        const float mul1 = (sPV[0] * mx + sPV[4] * my + sPV[8] * mz + sPV[12]) * m_w * m_w;
        const float mul2 = (sPV[1] * mx + sPV[5] * my + sPV[9] * mz + sPV[13]) * m_w * m_w;
        dmean[0] += (sPV[0] * m_w - sPV[3] * mul1) * g2x + (sPV[1] * m_w - sPV[3] * mul2) * g2y;
        dmean[1] += (sPV[4] * m_w - sPV[7] * mul1) * g2x + (sPV[5] * m_w - sPV[7] * mul2) * g2y;
        dmean[2] += (sPV[8] * m_w - sPV[11] * mul1) * g2x + (sPV[9] * m_w - sPV[11] * mul2) * g2y;
    }

    // ---- 4. SH adjoint incl. view-direction term     backward.cu:23-142 ----
    if (a.shs) {
        const float ox = mx - sCam[0], oy = my - sCam[1], oz = mz - sCam[2];
        const float sum2 = ox * ox + oy * oy + oz * oz;
        const float inv_len = 1.f / sqrtf(sum2);
        const float x = ox * inv_len, y = oy * inv_len, z = oz * inv_len;
        const uint8_t cl = a.clamped[i];
        const float gr = (cl & 1) ? 0.f : dcol[0], gg = (cl & 2) ? 0.f : dcol[1], gb = (cl & 4) ? 0.f : dcol[2];
        const float* sh = a.shs + (size_t)3 * a.M * i;
        float* dsh = a.g.dL_dsh ? a.g.dL_dsh + (size_t)3 * a.M * i : nullptr;
        float ddx = 0.f, ddy = 0.f, ddz = 0.f;   // dL/d(dir)
        int used = 1;
        // 16-byte aligned gradient rows (M = 4, 16): the basis values are kept and the row is written (or, in
        // accumulate mode, updated by read-modify-write) as float4 after the sweep: 12 instead of 48 accesses at degree 3
        const bool vecd = dsh != nullptr && (a.M & 3) == 0 && (reinterpret_cast<size_t>(a.g.dL_dsh) & 15u) == 0;
        float bv[16];
#pragma unroll
        for (int k = 0; k < 16; k++) bv[k] = 0.f;
        auto sh_section = [&](auto fetch) {
        // basis value b and its partials (bx,by,bz) per coefficient; dsh_k = b * g, ddir += partial * <sh_k, g>
        auto coef = [&](int k, float b, float bx, float by, float bz) {
            bv[k] = b;
            if (dsh && !vecd) { put<ACC>(dsh, 3 * k, b * gr); put<ACC>(dsh, 3 * k + 1, b * gg); put<ACC>(dsh, 3 * k + 2, b * gb); }
            const float sg = fetch(3 * k) * gr + fetch(3 * k + 1) * gg + fetch(3 * k + 2) * gb;
            ddx += bx * sg; ddy += by * sg; ddz += bz * sg;
        };
        coef(0, GS_SH_C0, 0.f, 0.f, 0.f);
        if (a.D > 0) {
            coef(1, -GS_SH_C1 * y, 0.f, -GS_SH_C1, 0.f);
            coef(2, GS_SH_C1 * z, 0.f, 0.f, GS_SH_C1);
            coef(3, -GS_SH_C1 * x, -GS_SH_C1, 0.f, 0.f);
            used = 4;
            if (a.D > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                coef(4, GS_SH_C2_0 * xy, GS_SH_C2_0 * y, GS_SH_C2_0 * x, 0.f);
                coef(5, GS_SH_C2_1 * yz, 0.f, GS_SH_C2_1 * z, GS_SH_C2_1 * y);
                coef(6, GS_SH_C2_2 * (2.f * zz - xx - yy), GS_SH_C2_2 * 2.f * -x, GS_SH_C2_2 * 2.f * -y, GS_SH_C2_2 * 2.f * 2.f * z);
                coef(7, GS_SH_C2_3 * xz, GS_SH_C2_3 * z, 0.f, GS_SH_C2_3 * x);
                coef(8, GS_SH_C2_4 * (xx - yy), GS_SH_C2_4 * 2.f * x, GS_SH_C2_4 * 2.f * -y, 0.f);
                used = 9;
                if (a.D > 2) {
                    coef(9, GS_SH_C3_0 * y * (3.f * xx - yy), GS_SH_C3_0 * 3.f * 2.f * xy, GS_SH_C3_0 * 3.f * (xx - yy), 0.f);
                    coef(10, GS_SH_C3_1 * xy * z, GS_SH_C3_1 * yz, GS_SH_C3_1 * xz, GS_SH_C3_1 * xy);
                    coef(11, GS_SH_C3_2 * y * (4.f * zz - xx - yy), GS_SH_C3_2 * -2.f * xy, GS_SH_C3_2 * (-3.f * yy + 4.f * zz - xx),
                         GS_SH_C3_2 * 4.f * 2.f * yz);
                    coef(12, GS_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy), GS_SH_C3_3 * -3.f * 2.f * xz, GS_SH_C3_3 * -3.f * 2.f * yz,
                         GS_SH_C3_3 * 3.f * (2.f * zz - xx - yy));
                    coef(13, GS_SH_C3_4 * x * (4.f * zz - xx - yy), GS_SH_C3_4 * (-3.f * xx + 4.f * zz - yy), GS_SH_C3_4 * -2.f * xy,
                         GS_SH_C3_4 * 4.f * 2.f * xz);
                    coef(14, GS_SH_C3_5 * z * (xx - yy), GS_SH_C3_5 * 2.f * xz, GS_SH_C3_5 * -2.f * yz, GS_SH_C3_5 * (xx - yy));
                    coef(15, GS_SH_C3_6 * x * (xx - 3.f * yy), GS_SH_C3_6 * 3.f * (xx - yy), GS_SH_C3_6 * -3.f * 2.f * xy, 0.f);
                    used = 16;
                }
            }
        }
        };
        if (sh_rows_aligned(a.shs, a.M)) sh_section([&](int f) { return sh_elem4(sh, f); });   // one LDG.128 per 4 floats
        else sh_section([&](int f) { return __ldg(sh + f); });
        if (vecd) {
            float4* d4 = reinterpret_cast<float4*>(dsh);
            const float g3[3] = {gr, gg, gb};
            // coefficients >= used have bv = 0: accumulate mode skips those chunks, overwrite mode stores zeros.
            // All loads of the row are issued before its first store: the sinks are not __restrict__, a load placed
            // after a store would wait for it (12 dependent round trips per Gaussian instead of one).
            auto update_row = [&]() {
                float4 q[12];
#pragma unroll
                for (int c = 0; c < 12; c++)
                    q[c] = (ACC && 4 * c < 3 * used && 4 * c < 3 * a.M) ? d4[c] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int c = 0; c < 12; c++) {
                    if ((!ACC || 4 * c < 3 * used) && 4 * c < 3 * a.M) {
                        q[c].x += bv[(4 * c) / 3] * g3[(4 * c) % 3];
                        q[c].y += bv[(4 * c + 1) / 3] * g3[(4 * c + 1) % 3];
                        q[c].z += bv[(4 * c + 2) / 3] * g3[(4 * c + 2) % 3];
                        q[c].w += bv[(4 * c + 3) / 3] * g3[(4 * c + 3) % 3];
                        d4[c] = q[c];
                    }
                }
            };
            if (PHASED) {
                for (int ph = 0; ph <= maxrank; ph++) {
                    if (rank == ph) update_row();
                    __syncwarp(syncmask);
                }
            } else {
                update_row();
            }
        }
        if (!ACC && dsh && !vecd)
            for (int k = 3 * used; k < 3 * a.M; k++) dsh[k] = 0.f;
        // dnormvdv, auxiliary.h:153-164
        const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
        dmean[0] += ((+sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * invsum32;
        dmean[1] += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * invsum32;
        dmean[2] += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * invsum32;
    }

    // ---- 5. cov3D -> scale / quaternion     backward.cu:330-393 ----
    float sr[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};     // d scale[3], d quaternion[4]
    if (a.scales) {
        const float4 q = reinterpret_cast<const float4*>(a.rotations)[i];
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        m3<float> R, Sm;
        R.m[0][0] = 1.f - 2.f * (y * y + z * z); R.m[0][1] = 2.f * (x * y - r * z); R.m[0][2] = 2.f * (x * z + r * y);
        R.m[1][0] = 2.f * (x * y + r * z); R.m[1][1] = 1.f - 2.f * (x * x + z * z); R.m[1][2] = 2.f * (y * z - r * x);
        R.m[2][0] = 2.f * (x * z - r * y); R.m[2][1] = 2.f * (y * z + r * x); R.m[2][2] = 1.f - 2.f * (x * x + y * y);
        const float s[3] = {a.scale_modifier * a.scales[3 * i], a.scale_modifier * a.scales[3 * i + 1],
                            a.scale_modifier * a.scales[3 * i + 2]};
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int rr = 0; rr < 3; rr++) Sm.m[c][rr] = (c == rr) ? s[c] : 0.f;
        const m3<float> M = mul(Sm, R);
        m3<float> dSig;
        dSig.m[0][0] = dcov[0]; dSig.m[0][1] = 0.5f * dcov[1]; dSig.m[0][2] = 0.5f * dcov[2];
        dSig.m[1][0] = 0.5f * dcov[1]; dSig.m[1][1] = dcov[3]; dSig.m[1][2] = 0.5f * dcov[4];
        dSig.m[2][0] = 0.5f * dcov[2]; dSig.m[2][1] = 0.5f * dcov[4]; dSig.m[2][2] = dcov[5];
        m3<float> dM = mul(M, dSig);
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int rr = 0; rr < 3; rr++) dM.m[c][rr] *= 2.0f;
        const m3<float> Rt = transpose(R);
        m3<float> dMt = transpose(dM);
        float dscale[3];
#pragma unroll
        for (int c = 0; c < 3; c++)
            dscale[c] = Rt.m[c][0] * dMt.m[c][0] + Rt.m[c][1] * dMt.m[c][1] + Rt.m[c][2] * dMt.m[c][2];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int rr = 0; rr < 3; rr++) dMt.m[c][rr] *= s[c];
        auto Dm = [&](int c, int rr) { return dMt.m[c][rr]; };
        const float dq0 = 2.f * z * (Dm(0, 1) - Dm(1, 0)) + 2.f * y * (Dm(2, 0) - Dm(0, 2)) + 2.f * x * (Dm(1, 2) - Dm(2, 1));
        const float dq1 = 2.f * y * (Dm(1, 0) + Dm(0, 1)) + 2.f * z * (Dm(2, 0) + Dm(0, 2)) + 2.f * r * (Dm(1, 2) - Dm(2, 1)) -
                          4.f * x * (Dm(2, 2) + Dm(1, 1));
        const float dq2 = 2.f * x * (Dm(1, 0) + Dm(0, 1)) + 2.f * r * (Dm(2, 0) - Dm(0, 2)) + 2.f * z * (Dm(1, 2) + Dm(2, 1)) -
                          4.f * y * (Dm(2, 2) + Dm(0, 0));
        const float dq3 = 2.f * r * (Dm(0, 1) - Dm(1, 0)) + 2.f * x * (Dm(2, 0) + Dm(0, 2)) + 2.f * y * (Dm(1, 2) + Dm(2, 1)) -
                          4.f * z * (Dm(1, 1) + Dm(0, 0));
        // the reference writes the gradient w.r.t. the *modified* scale (no scale_modifier factor), backward.cu:371-375
        for (int k = 0; k < 3; k++) sr[k] = dscale[k];
        sr[3] = dq0; sr[4] = dq1; sr[5] = dq2; sr[6] = dq3;
    }
    if (!a.shs && !ACC && a.g.dL_dsh)
        for (int k = 0; k < 3 * a.M; k++) a.g.dL_dsh[(size_t)3 * a.M * i + k] = 0.f;

    // ---- 6. the 22 small sinks: in accumulate mode every old value is loaded before the first store (the sinks are
    //         not __restrict__: interleaved read-modify-writes would be 22 dependent round trips) ----
    const bool w_sr = a.scales != nullptr || !ACC;     // without scales the overwrite mode still stores zeros
    float* const sc = w_sr ? a.g.dL_dscales : nullptr;
    float* const rq = w_sr ? a.g.dL_drotations : nullptr;
    auto update_small = [&]() {
        auto ld = [&](const float* p, size_t j) { return (ACC && p) ? p[j] : 0.f; };
        float o_sr[7], o_m2[2], o_col[3], o_m3[3], o_cov[6];
#pragma unroll
        for (int k = 0; k < 3; k++) o_sr[k] = ld(sc, 3 * i + k);
#pragma unroll
        for (int k = 0; k < 4; k++) o_sr[3 + k] = ld(rq, 4 * i + k);
        o_m2[0] = ld(a.g.dL_dmeans2D, 3 * i); o_m2[1] = ld(a.g.dL_dmeans2D, 3 * i + 1);
#pragma unroll
        for (int k = 0; k < 3; k++) { o_col[k] = ld(a.g.dL_dcolors, 3 * i + k); o_m3[k] = ld(a.g.dL_dmeans3D, 3 * i + k); }
        const float o_op = ld(a.g.dL_dopacity, i);
#pragma unroll
        for (int k = 0; k < 6; k++) o_cov[k] = ld(a.g.dL_dcov3D, 6 * i + k);
        auto st = [&](float* p, size_t j, float old, float v) { if (p) p[j] = ACC ? old + v : v; };
#pragma unroll
        for (int k = 0; k < 3; k++) st(sc, 3 * i + k, o_sr[k], sr[k]);
#pragma unroll
        for (int k = 0; k < 4; k++) st(rq, 4 * i + k, o_sr[3 + k], sr[3 + k]);
        st(a.g.dL_dmeans2D, 3 * i, o_m2[0], g2x);
        st(a.g.dL_dmeans2D, 3 * i + 1, o_m2[1], g2y);
        if (!ACC) st(a.g.dL_dmeans2D, 3 * i + 2, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < 3; k++) st(a.g.dL_dcolors, 3 * i + k, o_col[k], dcol[k]);
        st(a.g.dL_dopacity, i, o_op, dL_dop);
#pragma unroll
        for (int k = 0; k < 3; k++) st(a.g.dL_dmeans3D, 3 * i + k, o_m3[k], dmean[k]);
#pragma unroll
        for (int k = 0; k < 6; k++) st(a.g.dL_dcov3D, 6 * i + k, o_cov[k], dcov[k]);
    };
    if (PHASED) {
        for (int ph = 0; ph <= maxrank; ph++) {
            if (rank == ph) update_small();
            __syncwarp(syncmask);
        }
    } else {
        update_small();
    }
}

template <bool ACC>
__global__ void __launch_bounds__(128) k_preprocess_bwd(BwdArgs a) {
    if (ACC && blockIdx.x * blockDim.x >= *a.n_active) return;   // accumulate mode runs on the active list
    __shared__ float sV[16], sPV[16], sCam[3];
    if (threadIdx.x < 16) { sV[threadIdx.x] = a.viewmatrix[threadIdx.x]; sPV[threadIdx.x] = a.projmatrix[threadIdx.x]; }
    if (threadIdx.x < 3) sCam[threadIdx.x] = a.campos ? a.campos[threadIdx.x] : 0.f;
    __syncthreads();
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (ACC) {
        if ((uint32_t)idx >= *a.n_active) return;
        idx = (int)__ldg(a.active_ids + idx);
    } else if (idx >= a.P) return;
    bwd_gaussian<ACC>(a, sV, sPV, sCam, idx);
}

// ---------------------------------------------------------------------------------------------
// Multi-view variant (the CG path): ONE launch accumulates the adjoints of up to MAX_BATCH_VIEWS linearised views.
// A warp owns the Gaussians [ACT_RANGE r, ACT_RANGE (r+1)) for the whole launch and walks the views in order; in
// each view it visits that range's slice of the view's (ascending) active list, so
//   * no two warps ever touch the same output element: plain read-modify-write, no atomics, and the sum over views
//     is taken in view order -- deterministic;
//   * all warps are in flight at once and each hides the others' memory latency (the per-view launches this
//     replaces ran at 20 % occupancy behind a chain of dependent loads, 0.1 ms per view).
// ---------------------------------------------------------------------------------------------
struct BwdView {
    const float *viewmatrix, *projmatrix, *campos;
    const int32_t* radii;
    const uint8_t* clamped;
    const uint32_t* g_off;
    const float* cov3D;
    const float4* inst_grads;
    const uint32_t* active_ids;
    const uint32_t* act_range;
    int W, H;
    float tan_fovx, tan_fovy;
};
struct BwdBatch { BwdView v[MAX_BATCH_VIEWS]; int n; };

#ifndef GSLM_BWDV_MINB
#define GSLM_BWDV_MINB 4   // 128 registers: the per-Gaussian chain needs the ILP (96 registers: 1.52 ms per 16-view launch, 128: 0.95; profiles/r1i_adjoint_variants.txt)
#endif
__global__ void __launch_bounds__(128, GSLM_BWDV_MINB)
k_preprocess_bwd_views(BwdArgs a, const __grid_constant__ BwdBatch vb, int range_lo, int range_hi) {
    __shared__ float sCamAll[MAX_BATCH_VIEWS][36];   // V[16] | PV[16] | campos[3]
    for (int t = threadIdx.x; t < vb.n * 36; t += blockDim.x) {
        const int v = t / 36, k = t % 36;
        const BwdView& vw = vb.v[v];
        float x = 0.f;
        if (k < 16) x = vw.viewmatrix[k];
        else if (k < 32) x = vw.projmatrix[k - 16];
        else if (k < 35 && vw.campos) x = vw.campos[k - 32];
        sCamAll[v][k] = x;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int r = range_lo + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= range_hi) return;
    for (int v = 0; v < vb.n; v++) {
        const BwdView& vw = vb.v[v];
        const uint32_t lo = __ldg(vw.act_range + r), hi = __ldg(vw.act_range + r + 1);
        if (lo < hi) {
            BwdArgs b = a;
            b.W = vw.W; b.H = vw.H; b.tan_fovx = vw.tan_fovx; b.tan_fovy = vw.tan_fovy;
            b.radii = vw.radii; b.clamped = vw.clamped; b.g_off = vw.g_off; b.inst_grads = vw.inst_grads;
            if (vw.cov3D) b.cov3D = vw.cov3D;
            for (uint32_t s = lo + lane; s < hi; s += 32)
                bwd_gaussian<true>(b, sCamAll[v], sCamAll[v] + 16, sCamAll[v] + 32, (int)__ldg(vw.active_ids + s));
        }
        __syncwarp();   // the next view's read-modify-write of an element may come from another lane of this warp
    }
}

// ---------------------------------------------------------------------------------------------
// Dense variant (round 2).  The kernel above gives every view its own pass over the warp's slice of that view's active
// list: ~20 of the 256 Gaussians of a range are active per view, so a pass keeps ~20 of 32 lanes busy and the stage is
// 16 dependent-latency passes per launch.  Here the (view, active Gaussian) items of the range are ONE list in
// view-major order (prefix sums of the 16 slice lengths, per warp, in shared memory) and the lanes take 32 consecutive
// items per pass whatever view they belong to: ~10 full passes instead of 16 two-thirds-empty ones.  Two lanes of a pass
// can then hold the same Gaussian (for different views): __match_any_sync finds them and the PHASED store sites of
// bwd_gaussian let them update the row in view order, so the sums are taken in exactly the old order (deterministic;
// the compiler's fma contraction inside the two instantiations may differ in the last bit).
// ---------------------------------------------------------------------------------------------
#ifndef GSLM_BWDV_DENSE
#define GSLM_BWDV_DENSE 1
#endif
__global__ void __launch_bounds__(128, GSLM_BWDV_MINB)
k_preprocess_bwd_views_dense(BwdArgs a, const __grid_constant__ BwdBatch vb, int range_lo, int range_hi) {
    __shared__ float sCamAll[MAX_BATCH_VIEWS][36];   // V[16] | PV[16] | campos[3]
    __shared__ BwdView sView[MAX_BATCH_VIEWS];
    __shared__ uint32_t sLo[4][MAX_BATCH_VIEWS], sCum[4][MAX_BATCH_VIEWS + 1];
    for (int t = threadIdx.x; t < vb.n * 36; t += blockDim.x) {
        const int v = t / 36, k = t % 36;
        const BwdView& vw = vb.v[v];
        float x = 0.f;
        if (k < 16) x = vw.viewmatrix[k];
        else if (k < 32) x = vw.projmatrix[k - 16];
        else if (k < 35 && vw.campos) x = vw.campos[k - 32];
        sCamAll[v][k] = x;
    }
    if ((int)threadIdx.x < vb.n) sView[threadIdx.x] = vb.v[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int r = range_lo + blockIdx.x * (blockDim.x >> 5) + w;
    if (r >= range_hi) return;
    // the range's slice of every view's active list and the prefix sums of the slice lengths
    uint32_t lo = 0, cnt = 0;
    if (lane < vb.n) {
        lo = __ldg(sView[lane].act_range + r);
        cnt = __ldg(sView[lane].act_range + r + 1) - lo;
    }
    uint32_t inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    if (lane < vb.n) { sLo[w][lane] = lo; sCum[w][lane + 1] = inc; }
    if (lane == 0) sCum[w][0] = 0;
    const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
    __syncwarp();
    for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t t = base + lane;
        const bool valid = t < total;
        const uint32_t callmask = __ballot_sync(0xffffffffu, valid);
        if (valid) {
            int v = 0;
            while (v + 1 < vb.n && sCum[w][v + 1] <= t) v++;
            const BwdView& vw = sView[v];
            const int idx = (int)__ldg(vw.active_ids + sLo[w][v] + (t - sCum[w][v]));
            const uint32_t same = __match_any_sync(callmask, idx);
            const int rank = __popc(same & ((1u << lane) - 1u));
            const int maxrank = __reduce_max_sync(callmask, rank);
            BwdArgs b = a;
            b.W = vw.W; b.H = vw.H; b.tan_fovx = vw.tan_fovx; b.tan_fovy = vw.tan_fovy;
            b.radii = vw.radii; b.clamped = vw.clamped; b.g_off = vw.g_off; b.inst_grads = vw.inst_grads;
            if (vw.cov3D) b.cov3D = vw.cov3D;
            bwd_gaussian<true, true>(b, sCamAll[v], sCamAll[v] + 16, sCamAll[v] + 32, idx, callmask, rank, maxrank);
        }
        __syncwarp();   // the next pass may touch a row of this one from another lane
    }
}

static void fill_common(BwdArgs& a, const GslmGaussians& g, const GslmCamera& cam, bool has_invdepth_grad,
                        const GslmGrads& grads) {
    a.P = g.P; a.D = g.sh_degree; a.M = g.M; a.W = cam.width; a.H = cam.height;
    a.scale_modifier = g.scale_modifier; a.tan_fovx = cam.tan_fovx; a.tan_fovy = cam.tan_fovy;
    a.antialiasing = cam.antialiasing; a.has_invdepth = has_invdepth_grad ? 1 : 0;
    a.means3D = g.means3D; a.opacities = g.opacities; a.scales = g.cov3D_precomp ? nullptr : g.scales;
    a.rotations = g.cov3D_precomp ? nullptr : g.rotations;
    a.shs = g.colors_precomp ? nullptr : g.shs;
    a.viewmatrix = cam.viewmatrix; a.projmatrix = cam.projmatrix; a.campos = cam.campos;
    a.g = grads;
}

int launch_preprocess_bwd(const GslmGaussians& g, const GslmCamera& cam, const int32_t* radii, const GeomView& gv,
                          const BinningView& bv, const float4* inst_grads, bool has_invdepth_grad,
                          const GslmGrads& grads, cudaStream_t s) {
    if (g.P == 0) return 0;
    (void)bv;
    BwdArgs a;
    fill_common(a, g, cam, has_invdepth_grad, grads);
    a.cov3D = g.cov3D_precomp ? g.cov3D_precomp : gv.cov3D;
    a.radii = radii; a.clamped = gv.clamped; a.g_off = gv.g_off;
    a.active_ids = gv.active_ids; a.n_active = gv.n_active;
    a.inst_grads = inst_grads;
    int grid = (g.P + 127) / 128;
    StageScope ts(ST_PRE_BWD, s);
    if (grads.accumulate) k_preprocess_bwd<true><<<grid, 128, 0, s>>>(a);
    else k_preprocess_bwd<false><<<grid, 128, 0, s>>>(a);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_preprocess_bwd_views(const GslmGaussians& g, int n_views, const GslmCamera* cams, const int32_t* const* radii,
                                const GeomView* gvs, const float4* const* inst_grads, const GslmGrads& grads,
                                cudaStream_t s, int range_lo, int range_hi) {
    if (g.P == 0 || n_views == 0) return 0;
    if (n_views > MAX_BATCH_VIEWS) { set_error("too many views in one batch"); return 1; }
    BwdArgs a;
    fill_common(a, g, cams[0], false, grads);
    a.cov3D = g.cov3D_precomp; a.radii = nullptr; a.clamped = nullptr; a.g_off = nullptr;
    a.active_ids = nullptr; a.n_active = nullptr; a.inst_grads = nullptr;
    BwdBatch vb;
    memset(&vb, 0, sizeof(vb));
    vb.n = n_views;
    for (int i = 0; i < n_views; i++) {
        BwdView& v = vb.v[i];
        if (cams[i].antialiasing != cams[0].antialiasing) { set_error("views of one batch must agree on antialiasing"); return 1; }
        v.viewmatrix = cams[i].viewmatrix; v.projmatrix = cams[i].projmatrix; v.campos = cams[i].campos;
        v.radii = radii[i]; v.clamped = gvs[i].clamped; v.g_off = gvs[i].g_off;
        v.cov3D = g.cov3D_precomp ? nullptr : gvs[i].cov3D;
        v.inst_grads = inst_grads[i]; v.active_ids = gvs[i].active_ids; v.act_range = gvs[i].act_range;
        v.W = cams[i].width; v.H = cams[i].height; v.tan_fovx = cams[i].tan_fovx; v.tan_fovy = cams[i].tan_fovy;
    }
    const int nranges = (g.P + ACT_RANGE - 1) / ACT_RANGE;
    if (range_hi < 0 || range_hi > nranges) range_hi = nranges;
    if (range_lo < 0) range_lo = 0;
    if (range_lo >= range_hi) return 0;
    StageScope ts(ST_PRE_BWD, s);
    // the dense kernel's phased stores cover the float4 SH-row path only (M % 4 == 0, 16-byte aligned rows) -- every
    // degree-0 / degree-3 layout the solver produces; anything else takes the per-view passes
    const bool dense = GSLM_BWDV_DENSE &&
                       (!grads.dL_dsh || !a.shs || ((g.M & 3) == 0 && (reinterpret_cast<size_t>(grads.dL_dsh) & 15u) == 0));
    if (dense)
        k_preprocess_bwd_views_dense<<<(range_hi - range_lo + 3) / 4, 128, 0, s>>>(a, vb, range_lo, range_hi);
    else
        k_preprocess_bwd_views<<<(range_hi - range_lo + 3) / 4, 128, 0, s>>>(a, vb, range_lo, range_hi);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

}  // namespace gslm
