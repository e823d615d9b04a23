// lt_terrain.cu -- device-side builder of the tunnel tile table.
//
// Replaces the per-tile Python loop of the reference's terrain generator for terrain_type = single_path:
//   go1_gym/utils/tunnel.py:51-217        Terrain.__init__ / make_terrain / add_terrain_to_map
//   go1_gym/utils/tunnel_fn.py:3-21       vec_plane_from_points (four planes per wedge, min over faces, max over wedges)
//   go1_gym/utils/tunnel_fn.py:99-163     TerrainFunctions.single_path
//   legged_robot_trajectory_tracking.py:1827-1840   height_samples_by_row_col * vertical_scale -> float32
// and writes the product's deduplicated table tile_heights [tiles, 2, L, W] (layer 0 ceiling, layer 1 floor, metres)
// directly.  The random draws stay on the host (numpy's generator, same order as the reference: legged_tracking_b200/
// terrain.py); everything per cell runs here.  The reference computes in float64 with individually rounded numpy ops, so
// the kernel spells every operation with __dmul_rn / __dadd_rn / __dsub_rn / __ddiv_rn (no FMA contraction): the table is
// bit-identical to the reference's (tests/golden/terrain_single_path_*.npz).
//
// One CTA per (tile, layer): the first 8 threads derive the face planes of the (at most 2) wedges into shared memory, then
// all threads stream the L x W cells of the tile layer (coalesced float32 stores).  HBM-bound: 4 B written per cell.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lt_b200.h"

#define LT_TER_THREADS 256

struct lt_plane { double dc, ac, bc; };        // height(x, y) = dc - ac x - bc y

__global__ void __launch_bounds__(LT_TER_THREADS) lt_tunnel_tiles_kernel(const int32_t* __restrict__ n_pyr, const double* __restrict__ wedges,
                                                                         const int32_t* __restrict__ window, const double* __restrict__ xs,
                                                                         const double* __restrict__ ys, const double* __restrict__ consts,
                                                                         float* __restrict__ out, int tile_rows, int tile_cols,
                                                                         int sub_rows, int sub_cols) {
    __shared__ lt_plane s_plane[2][4];
    const int tl = blockIdx.x;                     // tile * 2 + layer
    const int layer = tl & 1;
    const int np = n_pyr[tl];
    if (threadIdx.x < 8 && (int)(threadIdx.x >> 2) < np) {
        const int q = threadIdx.x >> 2, f = threadIdx.x & 3;
        const double* w = wedges + ((size_t)tl * 2 + q) * 5;          // (mean_x, mean_y, mean_z, pw, pl)
        const double mx = w[0], my = w[1], mz = w[2], pw = w[3], pl = w[4];
        // base corners 0..3 of the wedge (tunnel_fn.py:136-141): (+pw,+pl) (-pw,+pl) (-pw,-pl) (+pw,-pl), z = 0; face f = (corner f, corner f+1, apex)
        const double cx[4] = {__dadd_rn(pw, mx), __dadd_rn(-pw, mx), __dadd_rn(-pw, mx), __dadd_rn(pw, mx)};
        const double cy[4] = {__dadd_rn(pl, my), __dadd_rn(pl, my), __dadd_rn(-pl, my), __dadd_rn(-pl, my)};
        const int i1 = f, i2 = (f + 1) & 3;
        const double v1x = __dsub_rn(mx, cx[i1]), v1y = __dsub_rn(my, cy[i1]), v1z = __dsub_rn(mz, 0.0);
        const double v2x = __dsub_rn(mx, cx[i2]), v2y = __dsub_rn(my, cy[i2]), v2z = __dsub_rn(mz, 0.0);
        // np.cross: each component is two separately rounded products and one subtraction
        const double a = __dsub_rn(__dmul_rn(v1y, v2z), __dmul_rn(v1z, v2y));
        const double b = __dsub_rn(__dmul_rn(v1z, v2x), __dmul_rn(v1x, v2z));
        const double c = __dsub_rn(__dmul_rn(v1x, v2y), __dmul_rn(v1y, v2x));
        const double d = __dadd_rn(__dadd_rn(__dmul_rn(a, mx), __dmul_rn(b, my)), __dmul_rn(c, mz));       // np.sum(cp * p3, -1)
        lt_plane P;
        P.dc = __ddiv_rn(d, c);
        P.ac = __ddiv_rn(a, c);
        P.bc = __ddiv_rn(b, c);
        s_plane[q][f] = P;
    }
    __syncthreads();
    const double vs = consts[0], ceil_over_vs = consts[1], min_top = consts[2];
    const double border = layer == 0 ? consts[3] : consts[4];
    const float border_f = __double2float_rn(__dmul_rn(border, vs));
    const int r0 = window[2 * (tl >> 1)], c0 = window[2 * (tl >> 1) + 1];
    float* o = out + (size_t)tl * tile_rows * tile_cols;
    // one cell of the patch: u along the tunnel (py of the sub-terrain), v across (px)
    auto cell = [&](int u, int v) -> float {
        double h;
        if (layer == 1 && (u == 0 || u == sub_rows - 1 || v == 0 || v == sub_cols - 1)) {
            h = 0.5;                                             // tunnel_fn.py:155-159: the floor patch is walled in
        } else {
            const double x = xs[u], y = ys[v];
            h = 0.0;
            for (int q = 0; q < np; ++q) {
                double m = __longlong_as_double(0x7ff0000000000000ll);
#pragma unroll
                for (int f = 0; f < 4; ++f) {
                    const lt_plane P = s_plane[q][f];
                    double t = __dsub_rn(__dsub_rn(P.dc, __dmul_rn(P.ac, x)), __dmul_rn(P.bc, y));
                    t = t < 0.0 ? 0.0 : t;                       // np.clip(., 0, inf)
                    m = t < m ? t : m;
                }
                h = (q == 0 || m > h) ? m : h;
            }
        }
        const double hi = (double)(long long)__ddiv_rn(h, vs);                // (h / vertical_scale).astype(int)
        double val = hi;
        if (layer == 0) {
            const double t = __dsub_rn(ceil_over_vs, hi);                     // tunnel.py:95: flip at the ceiling
            val = t < min_top ? min_top : t;                                  // tunnel.py:97: keep 0.05 m of clearance
        }
        return __double2float_rn(__dmul_rn(val, vs));                         // LR:1838-1840
    };
    if ((tile_cols & 3) == 0) {
        // 16-byte stores: a thread owns chunks of 4 consecutive cells of a row; three quarters of a tile are border constants
        const int cols4 = tile_cols >> 2, chunks = tile_rows * cols4;
        float4* o4 = reinterpret_cast<float4*>(o);
        for (int q = threadIdx.x; q < chunks; q += LT_TER_THREADS) {
            const int r = q / cols4, cc = (q - r * cols4) << 2;
            const int u = r - r0, v = cc - c0;
            float e[4] = {border_f, border_f, border_f, border_f};
            if (u >= 0 && u < sub_rows && v + 3 >= 0 && v < sub_cols) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (v + k >= 0 && v + k < sub_cols) e[k] = cell(u, v + k);
            }
            const float4 w = make_float4(e[0], e[1], e[2], e[3]);
            o4[q] = w;
        }
    } else {
        const int cells = tile_rows * tile_cols;
        for (int k = threadIdx.x; k < cells; k += LT_TER_THREADS) {
            const int r = k / tile_cols, cc = k - r * tile_cols;
            const int u = r - r0, v = cc - c0;
            o[k] = (u >= 0 && u < sub_rows && v >= 0 && v < sub_cols) ? cell(u, v) : border_f;
        }
    }
}

extern "C" int lt_build_tunnel_tiles(const int32_t* n_pyramids, const double* wedges, const int32_t* window, const double* xs,
                                     const double* ys, const double* consts, float* tile_heights, int32_t n_tiles, int32_t tile_rows,
                                     int32_t tile_cols, int32_t sub_rows, int32_t sub_cols, void* stream) {
    if (!n_pyramids || !wedges || !window || !xs || !ys || !consts || !tile_heights) return LT_ERR_INVALID;
    if (n_tiles <= 0 || tile_rows <= 0 || tile_cols <= 0 || sub_rows <= 0 || sub_cols <= 0) return LT_ERR_INVALID;
    lt_tunnel_tiles_kernel<<<(unsigned)n_tiles * 2u, LT_TER_THREADS, 0, (cudaStream_t)stream>>>(
        n_pyramids, wedges, window, xs, ys, consts, tile_heights, tile_rows, tile_cols, sub_rows, sub_cols);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}
