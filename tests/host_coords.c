/* Test infrastructure: the product's coordinate header (vfdepth_b200/csrc/vfd_coords.h) compiled as plain C
 * (gcc -O2 -ffp-contract=off) so that the exact arithmetic the CUDA kernels run can be compared bit for bit
 * with the oracle on a machine without a GPU.  Built and driven by tests/test_host_coords.py. */
#include "../vfdepth_b200/csrc/vfd_coords.h"

/* one camera: N voxels -> v_local [3][N] (matmul layout), grid [N][2] */
void hc_voxel_grid(const float* Einv, const float* K, const float* ax, const float* ay, const float* az,
                   int X, int Y, int Z, int w, int h, float* v_out, float* g_out) {
    long N = (long)X * Y * Z;
    for (long n = 0; n < N; ++n) {
        int xi = (int)(n % X), yi = (int)((n / X) % Y), zi = (int)(n / ((long)X * Y));
        float v[3], g[2];
        vfd_voxel_to_pixel(Einv, K, ax[xi], ay[yi], az[zi], w, h, v, g);
        v_out[n] = v[0]; v_out[N + n] = v[1]; v_out[2 * N + n] = v[2];
        g_out[2 * n] = g[0]; g_out[2 * n + 1] = g[1];
    }
}

/* one camera: grid [D][h][w][3] */
void hc_ray_grid(const float* invK, const float* E, const float* bins, int D, int h, int w,
                 const float* start, const float* length, float* g_out) {
    for (int d = 0; d < D; ++d)
        for (int v = 0; v < h; ++v)
            for (int u = 0; u < w; ++u) {
                float ray[3];
                vfd_pixel_ray(invK, (float)u, (float)v, ray);
                vfd_ray_point_to_grid(E, ray, bins[d], start, length, g_out + 3 * (((long)d * h + v) * w + u));
            }
}

/* one sample: grid [H][W][2] */
void hc_warp_grid(const float* invK, const float* K, const float* T, const float* depth, int H, int W, float* g_out) {
    float P[12];
    vfd_premul_KT(K, T, P);
    for (int v = 0; v < H; ++v)
        for (int u = 0; u < W; ++u) {
            float X[3], cam[3];
            vfd_backproject(invK, (float)u, (float)v, depth[(long)v * W + u], X);
            vfd_reproject(P, X, W, H, cam, g_out + 2 * ((long)v * W + u));
        }
}
