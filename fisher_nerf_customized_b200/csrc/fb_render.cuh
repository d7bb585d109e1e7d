// fb_render.cuh -- markVisible (the drop-in forward render itself is k_render_warp, fb_render_warp.cuh).
#pragma once
#include "fb_layout.cuh"

namespace fb {

// present[i] = (view * p_i).z > 0.001   (checkFrustum, rasterizer_impl.cu:54-66; the projection
// is computed by the reference but does not influence the result).
__global__ void k_mark_visible(int P, const float* __restrict__ means, Cam cam, uint8_t* __restrict__ present) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    float3 p = make_float3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
    float3 pv = xform4x3(p, cam.view);
    present[idx] = pv.z <= 0.001f ? 0 : 1;
}

}  // namespace fb
