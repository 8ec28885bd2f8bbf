// Coordinate arithmetic of the VFDepth hot path, written once for device and host.
//
// "Bit-exact voxel-to-pixel indices and validity masks" is only reachable by replaying the
// reference's fp32 operation ORDER.  The reference computes coordinates with torch.matmul
// (network/volumetric_fusionnet.py:133,170,245,251; models/geometry/geometry_util.py:58,67)
// followed by separate elementwise ops.  Measured in the build container: the CPU sgemm behind
// those matmuls with a contraction length of 3 or 4 and a long N is bit-identical to a
// sequential FMA chain  acc = a0*x0; acc = fma(a1,x1,acc); ...  while the small [4,4]@[4,4]
// product of geometry_util.py:67 (K @ T) is bit-identical to a sequential chain WITHOUT
// fused multiply-add.  Everything else is one rounding per torch op.
//
// On the device every operation uses the round-to-nearest intrinsics so that nvcc can neither
// contract nor reassociate; on the host (tests/host_coords.c, compiled with -ffp-contract=off)
// the same source runs as plain C and is checked against the oracle on the CPU.
#ifndef VFD_COORDS_H_
#define VFD_COORDS_H_

#include <math.h>

#if defined(__CUDACC__)
#define VFD_HD __host__ __device__ __forceinline__
#else
#define VFD_HD static inline
#endif

#if defined(__CUDA_ARCH__)
#define VFD_MUL(a, b) __fmul_rn((a), (b))
#define VFD_ADD(a, b) __fadd_rn((a), (b))
#define VFD_SUB(a, b) __fsub_rn((a), (b))
#define VFD_DIV(a, b) __fdiv_rn((a), (b))
#define VFD_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#else
#define VFD_MUL(a, b) ((a) * (b))
#define VFD_ADD(a, b) ((a) + (b))
#define VFD_SUB(a, b) ((a) - (b))
#define VFD_DIV(a, b) ((a) / (b))
#define VFD_FMA(a, b, c) fmaf((a), (b), (c))
#endif

// row . (x0,x1,x2,x3) as the long-N sgemm does it
VFD_HD float vfd_dot4(const float* r, float x0, float x1, float x2, float x3) {
    float acc = VFD_MUL(r[0], x0);
    acc = VFD_FMA(r[1], x1, acc);
    acc = VFD_FMA(r[2], x2, acc);
    acc = VFD_FMA(r[3], x3, acc);
    return acc;
}

VFD_HD float vfd_dot3(const float* r, float x0, float x1, float x2) {
    float acc = VFD_MUL(r[0], x0);
    acc = VFD_FMA(r[1], x1, acc);
    acc = VFD_FMA(r[2], x2, acc);
    return acc;
}

// row . column of a small [4,4]@[4,4] product (no FMA): geometry_util.py:67  (K @ T)
VFD_HD float vfd_dot4_nofma(const float* r, float x0, float x1, float x2, float x3) {
    float acc = VFD_MUL(r[0], x0);
    acc = VFD_ADD(acc, VFD_MUL(r[1], x1));
    acc = VFD_ADD(acc, VFD_MUL(r[2], x2));
    acc = VFD_ADD(acc, VFD_MUL(r[3], x3));
    return acc;
}

// ATen grid_sampler_unnormalize, align_corners=True: ((g + 1) / 2) * (size - 1).  x / 2 and x * 0.5 are the same
// correctly rounded value for every float, so the division is written as a multiplication (one instruction instead of ~9).
VFD_HD float vfd_unnormalize(float g, int size) {
    return VFD_MUL(VFD_MUL(VFD_ADD(g, 1.0f), 0.5f), (float)(size - 1));
}

// ---------------------------------------------------------------------------------------------
// voxel centre -> camera pixel   (volumetric_fusionnet.py:132-133, 166-181)
//   Einv : rows 0..2 of the 4x4 ego->camera matrix (12 floats, row stride 4)
//   K    : 4x4 intrinsics at the feature scale (uses [:3,:3], row stride 4)
// out: v[3] = point in the camera frame, g[2] = normalised sampling coordinate.
// The batch-global clamp of :173-174 only fires when some coordinate is non-finite and cannot
// change a validity decision (inf -> +-2w is still out of bounds, NaN stays NaN); not replayed.
// ---------------------------------------------------------------------------------------------
VFD_HD void vfd_voxel_to_pixel(const float* Einv, const float* K, float x, float y, float z, int w, int h,
                               float* v, float* g) {
    v[0] = vfd_dot4(Einv + 0, x, y, z, 1.0f);
    v[1] = vfd_dot4(Einv + 4, x, y, z, 1.0f);
    v[2] = vfd_dot4(Einv + 8, x, y, z, 1.0f);
    float cx = vfd_dot3(K + 0, v[0], v[1], v[2]);
    float cy = vfd_dot3(K + 4, v[0], v[1], v[2]);
    float cz = vfd_dot3(K + 8, v[0], v[1], v[2]);
    float den = VFD_ADD(cz, 1e-8f);
    float px = VFD_DIV(cx, den);
    float py = VFD_DIV(cy, den);
    px = VFD_DIV(px, (float)(w - 1));
    py = VFD_DIV(py, (float)(h - 1));
    g[0] = VFD_MUL(VFD_SUB(px, 0.5f), 2.0f);
    g[1] = VFD_MUL(VFD_SUB(py, 0.5f), 2.0f);
}

// strict out-of-bounds test of volumetric_fusionnet.py:192-193 / view_rendering.py:77-79:
// exactly +-1 is inside, NaN is inside (both comparisons false).
VFD_HD int vfd_in_bounds2(float gx, float gy) {
    return !((gx > 1.0f) || (gx < -1.0f) || (gy > 1.0f) || (gy < -1.0f));
}

// ---------------------------------------------------------------------------------------------
// feature pixel (u,v) at depth bin value `dep` -> normalised voxel-grid coordinate
// (volumetric_fusionnet.py:245-258).  ray[3] = invK[:3,:3] @ (u,v,1) is shared by all bins.
//   E     : rows 0..2 of the 4x4 camera->ego matrix
//   start : voxel_str_p[i] as fp32;  length : (voxel_end_p[i] - voxel_str_p[i]) as fp32
// ---------------------------------------------------------------------------------------------
VFD_HD void vfd_pixel_ray(const float* invK, float u, float v, float* ray) {
    ray[0] = vfd_dot3(invK + 0, u, v, 1.0f);
    ray[1] = vfd_dot3(invK + 4, u, v, 1.0f);
    ray[2] = vfd_dot3(invK + 8, u, v, 1.0f);
}

VFD_HD void vfd_ray_point_to_grid(const float* E, const float* ray, float dep, const float* start,
                                  const float* length, float* g) {
    float cx = VFD_MUL(dep, ray[0]);
    float cy = VFD_MUL(dep, ray[1]);
    float cz = VFD_MUL(dep, ray[2]);
    for (int i = 0; i < 3; ++i) {
        float p = vfd_dot4(E + 4 * i, cx, cy, cz, 1.0f);
        g[i] = VFD_SUB(VFD_MUL(VFD_DIV(VFD_SUB(p, start[i]), length[i]), 2.0f), 1.0f);
    }
}

// ---------------------------------------------------------------------------------------------
// target pixel (u,v) with depth `dep` -> normalised coordinate in the source image
// (geometry_util.py:52-77).  P = (K_src @ T)[:3,:] must come from vfd_premul_KT.
// ---------------------------------------------------------------------------------------------
VFD_HD void vfd_premul_KT(const float* K, const float* T, float* P /*12*/) {
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c)
            P[4 * r + c] = vfd_dot4_nofma(K + 4 * r, T[c], T[4 + c], T[8 + c], T[12 + c]);
}

VFD_HD void vfd_backproject(const float* invK, float u, float v, float dep, float* X) {
    float ray[3];
    vfd_pixel_ray(invK, u, v, ray);
    X[0] = VFD_MUL(dep, ray[0]);
    X[1] = VFD_MUL(dep, ray[1]);
    X[2] = VFD_MUL(dep, ray[2]);
}

VFD_HD void vfd_reproject(const float* P, const float* X, int W, int H, float* cam /*3*/, float* g /*2*/) {
    cam[0] = vfd_dot4(P + 0, X[0], X[1], X[2], 1.0f);
    cam[1] = vfd_dot4(P + 4, X[0], X[1], X[2], 1.0f);
    cam[2] = vfd_dot4(P + 8, X[0], X[1], X[2], 1.0f);
    float den = VFD_ADD(cam[2], 1e-7f);
    float px = VFD_DIV(VFD_DIV(cam[0], den), (float)(W - 1));
    float py = VFD_DIV(VFD_DIV(cam[1], den), (float)(H - 1));
    g[0] = VFD_MUL(VFD_SUB(px, 0.5f), 2.0f);
    g[1] = VFD_MUL(VFD_SUB(py, 0.5f), 2.0f);
}

#endif  // VFD_COORDS_H_
