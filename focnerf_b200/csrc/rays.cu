// rays.cu -- the batch a train step consumes, built on the device from what the host actually chooses: pixel indices.
//
// The reference generates rays on the device as well, but from ~15 torch launches per batch (nerf/utils.py:56-157, legacy/nerf/utils.py
// get_rays: meshgrid, gather, stack, norm, a [B,N,3]@[B,3,3] matmul) and its loader gathers the ground truth with an indexing kernel
// (nerf/provider.py collate: images[index, inds]; train_step: rgb * a + bg * (1 - a) for RGBA, legacy/nerf/utils.py:468-500).
// k_pixel_batch is both in one launch: ray n looks at pixel pix[n] of camera pose_idx[n]
//     i = (pix % W) + 0.5, j = (pix / W) + 0.5;  dir = ((i - cx) / fx, (j - cy) / fy, 1) / |.|;  rays_d = R dir;  rays_o = t
// and, when images are resident, gt[n] = images[pose_idx[n]][pix[n]] (RGB, or RGBA blended over bg).  A step then uploads 8 bytes per
// ray instead of 36.
#include "common.cuh"
#include "../../include/foc_b200.h"

using namespace foc;

namespace {

__global__ void __launch_bounds__(256) k_pixel_batch(const float* __restrict__ poses, uint32_t n_poses, float fx, float fy, float cx, float cy, uint32_t H, uint32_t W,
                                                     const int32_t* __restrict__ pose_idx, const int32_t* __restrict__ pix, uint32_t N,
                                                     float* __restrict__ rays_o, float* __restrict__ rays_d,
                                                     const float* __restrict__ images, uint32_t C, float bg, float* __restrict__ gt) {
    pdl_sync();
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    uint32_t p = pose_idx ? (uint32_t)__ldg(pose_idx + n) : 0u;
    if (p >= n_poses) p = n_poses - 1;
    uint32_t q = pix ? (uint32_t)__ldg(pix + n) : n;                   // no index list: pixel n of a full frame
    if (q >= H * W) q = H * W - 1;
    const float* M = poses + 16 * (size_t)p;                           // [4,4] row-major camera-to-world
    // explicit roundings in the order of the reference's separate torch ops (nerf/utils.py:125-131)
    const float i = __fadd_rn((float)(q % W), 0.5f), j = __fadd_rn((float)(q / W), 0.5f);
    const float xs = __fdiv_rn(__fsub_rn(i, cx), fx), ys = __fdiv_rn(__fsub_rn(j, cy), fy);
    const float nrm = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(xs, xs), __fmul_rn(ys, ys)), 1.0f));
    const float dx = __fdiv_rn(xs, nrm), dy = __fdiv_rn(ys, nrm), dz = __fdiv_rn(1.0f, nrm);
    #pragma unroll
    for (int k = 0; k < 3; ++k) {
        rays_d[3 * (size_t)n + k] = fmaf(dz, __ldg(M + 4 * k + 2), fmaf(dy, __ldg(M + 4 * k + 1), __fmul_rn(dx, __ldg(M + 4 * k))));
        rays_o[3 * (size_t)n + k] = __ldg(M + 4 * k + 3);
    }
    if (gt != nullptr) {
        const float* px = images + ((size_t)p * H * W + q) * C;
        if (C == 4) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(px));
            const float rest = __fmul_rn(bg, __fsub_rn(1.0f, v.w));
            gt[3 * (size_t)n] = __fadd_rn(__fmul_rn(v.x, v.w), rest); gt[3 * (size_t)n + 1] = __fadd_rn(__fmul_rn(v.y, v.w), rest);
            gt[3 * (size_t)n + 2] = __fadd_rn(__fmul_rn(v.z, v.w), rest);
        } else {
            gt[3 * (size_t)n] = __ldg(px); gt[3 * (size_t)n + 1] = __ldg(px + 1); gt[3 * (size_t)n + 2] = __ldg(px + 2);
        }
    }
}

}  // namespace

FOC_API int foc_pixel_batch(const float* poses, uint32_t n_poses, float fx, float fy, float cx, float cy, uint32_t H, uint32_t W,
                            const int32_t* pose_idx, const int32_t* pix, uint32_t N, float* rays_o, float* rays_d,
                            const float* images, uint32_t channels, float bg, float* gt, void* stream) {
    if (N == 0) return 0;
    if (poses == nullptr || n_poses == 0 || H == 0 || W == 0 || (uint64_t)H * W > 0x7fffffffull || rays_o == nullptr || rays_d == nullptr) return FOC_ERR_UNSUPPORTED;
    if (gt != nullptr && (images == nullptr || (channels != 3 && channels != 4))) return FOC_ERR_UNSUPPORTED;
    if (pix == nullptr && N > H * W) return FOC_ERR_UNSUPPORTED;
    foc::launch(k_pixel_batch, ceil_div(N, 256u), 256, 0, reinterpret_cast<cudaStream_t>(stream), poses, n_poses, fx, fy, cx, cy, H, W, pose_idx, pix, N,
                rays_o, rays_d, images, channels, bg, gt);
    FOC_LAUNCH_CHECK(); return 0;
}
