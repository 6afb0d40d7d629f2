// Host-model glue, NOT a row of SURVEY.md §8: the element-wise parts of the update block
// (MotionEncoder tail model.py:286-290, ConvGRU model.py:293-312) fused around the library
// convolutions.  Tensors are NHWC; the hidden state, the context and the motion features live in
// channel slices of the two GRU convolution input buffers, so the reference's concatenations vanish:
//
//   buffer A (z/r conv input) = [ h      | context | motion features | flow ]
//   buffer B (q conv input)   = [ r * h  | context | motion features | flow ]
//
//   motion_scatter: relu(cf_pre + bias) and flow -> the motion slots of A and B
//   gate:    z  = sigmoid(zr_pre[:, :H] + bz);  B.h-slot = sigmoid(zr_pre[:, H:] + br) * A.h-slot
//   update:  A.h-slot = (1 - z) * h + z * tanh(q_pre + bq)   (also to a dense copy for the heads)
// zr_pre is the output of ONE convolution with the z and r filters stacked (same products, one launch).

#include "raftk_common.cuh"

namespace raftk {

template <int V>
struct alignas(4 * V) Vec {
  float v[V];
};

__device__ __forceinline__ float sigmoidf_(float v) { return 1.0f / (1.0f + expf(-v)); }

template <int V>
__global__ void __launch_bounds__(256)
gru_gate_kernel(const float* __restrict__ zr_pre, const float* __restrict__ bias_zr, const float* __restrict__ h, int ldh,
                float* __restrict__ z_out, float* __restrict__ rh_out, int ld_rh, long long npix, int H) {
  pdl_wait();
  using T = Vec<V>;
  const int HV = H / V;
  const long long total = npix * HV;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long pix = t / HV;
    const int c = (int)(t - pix * HV) * V;
    const T zp = *reinterpret_cast<const T*>(zr_pre + pix * 2 * H + c);
    const T rp = *reinterpret_cast<const T*>(zr_pre + pix * 2 * H + H + c);
    const T bz = *reinterpret_cast<const T*>(bias_zr + c);
    const T br = *reinterpret_cast<const T*>(bias_zr + H + c);
    const T hv = *reinterpret_cast<const T*>(h + pix * ldh + c);
    T z, rh;
#pragma unroll
    for (int k = 0; k < V; ++k) {
      z.v[k] = sigmoidf_(zp.v[k] + bz.v[k]);
      rh.v[k] = sigmoidf_(rp.v[k] + br.v[k]) * hv.v[k];
    }
    *reinterpret_cast<T*>(z_out + pix * H + c) = z;
    *reinterpret_cast<T*>(rh_out + pix * ld_rh + c) = rh;
  }
}

template <int V>
__global__ void __launch_bounds__(256)
gru_update_kernel(const float* __restrict__ q_pre, const float* __restrict__ bq, const float* __restrict__ z,
                  float* __restrict__ h, int ldh, float* __restrict__ h_dense, long long npix, int H) {
  pdl_wait();
  using T = Vec<V>;
  const int HV = H / V;
  const long long total = npix * HV;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long pix = t / HV;
    const int c = (int)(t - pix * HV) * V;
    const T qp = *reinterpret_cast<const T*>(q_pre + pix * H + c);
    const T b = *reinterpret_cast<const T*>(bq + c);
    const T zv = *reinterpret_cast<const T*>(z + pix * H + c);
    T hv = *reinterpret_cast<const T*>(h + pix * ldh + c);
#pragma unroll
    for (int k = 0; k < V; ++k) hv.v[k] = (1.0f - zv.v[k]) * hv.v[k] + zv.v[k] * tanhf(qp.v[k] + b.v[k]);
    *reinterpret_cast<T*>(h + pix * ldh + c) = hv;
    *reinterpret_cast<T*>(h_dense + pix * H + c) = hv;
  }
}

// c_cf conv channels (bias + relu) followed by the 2 flow channels, into two strided destinations.
__global__ void __launch_bounds__(256)
motion_scatter_kernel(const float* __restrict__ cf_pre, const float* __restrict__ bias, const float* __restrict__ flow,
                      float* __restrict__ dst_a, float* __restrict__ dst_b, int ld, long long npix, int c_cf, int ld_src) {
  pdl_wait();
  const int per_pix = c_cf / 2 + 1;
  const long long total = npix * per_pix;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long pix = t / per_pix;
    const int c = (int)(t - pix * per_pix) * 2;
    float2 v;
    if (c < c_cf) {
      const float2 p = *reinterpret_cast<const float2*>(cf_pre + pix * ld_src + c);
      const float2 b = *reinterpret_cast<const float2*>(bias + c);
      v.x = fmaxf(p.x + b.x, 0.0f);
      v.y = fmaxf(p.y + b.y, 0.0f);
    } else {
      v = *reinterpret_cast<const float2*>(flow + pix * 2);
    }
    *reinterpret_cast<float2*>(dst_a + pix * ld + c) = v;
    *reinterpret_cast<float2*>(dst_b + pix * ld + c) = v;
  }
}

// UpdateCell tail (model.py:504-506): coords1 += delta_pre + bias;  flow = coords1 - coords0.
__global__ void __launch_bounds__(256)
flow_update_kernel(const float* __restrict__ delta_pre, int ld_delta, const float2* __restrict__ bias_p,
                   const float2* __restrict__ coords0, float2* __restrict__ coords1, float2* __restrict__ flow,
                   float* __restrict__ flow_pitched, int ld_flow, long long npix) {
  pdl_wait();
  // the kernel after this one in the step is the lookup, which waits for this grid in its first instruction:
  // let it be scheduled now
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const float2 bias = __ldg(bias_p);
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < npix; t += (long long)gridDim.x * blockDim.x) {
    const float2 d = *reinterpret_cast<const float2*>(delta_pre + t * ld_delta), c0 = coords0[t];
    float2 c1 = coords1[t];
    c1.x = __fadd_rn(c1.x, __fadd_rn(d.x, bias.x));
    c1.y = __fadd_rn(c1.y, __fadd_rn(d.y, bias.y));
    coords1[t] = c1;
    const float2 f = make_float2(__fadd_rn(c1.x, -c0.x), __fadd_rn(c1.y, -c0.y));
    flow[t] = f;
    if (flow_pitched) *reinterpret_cast<float2*>(flow_pitched + t * ld_flow) = f;
  }
}

static int grid_for(long long total) {
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)RAFTK_SM_COUNT * 16) blocks = (long long)RAFTK_SM_COUNT * 16;
  return (int)blocks;
}

}  // namespace raftk

extern "C" int raftk_gru_gate(const float* zr_pre, const float* bias_zr, const float* h, int ldh, float* z_out,
                              float* rh_out, int ld_rh, long long npix, int hidden, void* stream) {
  using namespace raftk;
  if (npix < 0 || hidden <= 0 || hidden % 2 || ldh % 2 || ld_rh % 2 || ldh < hidden || ld_rh < hidden) return RAFTK_ERR_SHAPE;
  if (npix == 0) return RAFTK_OK;
  if (!zr_pre || !bias_zr || !h || !z_out || !rh_out) return RAFTK_ERR_NULL;
  const uintptr_t bits = (uintptr_t)zr_pre | (uintptr_t)bias_zr | (uintptr_t)h | (uintptr_t)z_out | (uintptr_t)rh_out;
  if (bits & 7) return RAFTK_ERR_ALIGN;
  const bool v4 = !(bits & 15) && hidden % 4 == 0 && ldh % 4 == 0 && ld_rh % 4 == 0;
  auto s = static_cast<cudaStream_t>(stream);
  if (v4) return launch_pdl(gru_gate_kernel<4>, grid_for(npix * (hidden / 4)), 256, 0, s, zr_pre, bias_zr, h, ldh, z_out, rh_out, ld_rh, npix, hidden);
  return launch_pdl(gru_gate_kernel<2>, grid_for(npix * (hidden / 2)), 256, 0, s, zr_pre, bias_zr, h, ldh, z_out, rh_out, ld_rh, npix, hidden);
}

extern "C" int raftk_gru_update(const float* q_pre, const float* bq, const float* z, float* h, int ldh, float* h_dense,
                                long long npix, int hidden, void* stream) {
  using namespace raftk;
  if (npix < 0 || hidden <= 0 || hidden % 2 || ldh % 2 || ldh < hidden) return RAFTK_ERR_SHAPE;
  if (npix == 0) return RAFTK_OK;
  if (!q_pre || !bq || !z || !h || !h_dense) return RAFTK_ERR_NULL;
  const uintptr_t bits = (uintptr_t)q_pre | (uintptr_t)bq | (uintptr_t)z | (uintptr_t)h | (uintptr_t)h_dense;
  if (bits & 7) return RAFTK_ERR_ALIGN;
  const bool v4 = !(bits & 15) && hidden % 4 == 0 && ldh % 4 == 0;
  auto s = static_cast<cudaStream_t>(stream);
  if (v4) return launch_pdl(gru_update_kernel<4>, grid_for(npix * (hidden / 4)), 256, 0, s, q_pre, bq, z, h, ldh, h_dense, npix, hidden);
  return launch_pdl(gru_update_kernel<2>, grid_for(npix * (hidden / 2)), 256, 0, s, q_pre, bq, z, h, ldh, h_dense, npix, hidden);
}

extern "C" int raftk_motion_scatter(const float* cf_pre, int ld_src, const float* bias, const float* flow, float* dst_a,
                                    float* dst_b, int ld, long long npix, int c_cf, void* stream) {
  using namespace raftk;
  if (npix < 0 || c_cf <= 0 || c_cf % 2 || ld % 2 || ld < c_cf + 2 || ld_src % 2 || ld_src < c_cf) return RAFTK_ERR_SHAPE;
  if (npix == 0) return RAFTK_OK;
  if (!cf_pre || !bias || !flow || !dst_a || !dst_b) return RAFTK_ERR_NULL;
  if (((uintptr_t)cf_pre | (uintptr_t)bias | (uintptr_t)flow | (uintptr_t)dst_a | (uintptr_t)dst_b) & 7) return RAFTK_ERR_ALIGN;
  return launch_pdl(motion_scatter_kernel, grid_for(npix * (c_cf / 2 + 1)), 256, 0, static_cast<cudaStream_t>(stream), cf_pre, bias,
             flow, dst_a, dst_b, ld, npix, c_cf, ld_src);
}

extern "C" int raftk_flow_update(const float* delta_pre, int ld_delta, const float* bias, const float* coords0,
                                 float* coords1, float* flow, float* flow_pitched, int ld_flow, long long npix,
                                 void* stream) {
  using namespace raftk;
  if (npix < 0 || ld_delta < 2 || ld_delta % 2 || (flow_pitched && (ld_flow < 2 || ld_flow % 2))) return RAFTK_ERR_SHAPE;
  if (npix == 0) return RAFTK_OK;
  if (!delta_pre || !bias || !coords0 || !coords1 || !flow) return RAFTK_ERR_NULL;
  if (((uintptr_t)delta_pre | (uintptr_t)bias | (uintptr_t)coords0 | (uintptr_t)coords1 | (uintptr_t)flow |
       (uintptr_t)flow_pitched) & 7)
    return RAFTK_ERR_ALIGN;
  return launch_pdl(flow_update_kernel, grid_for(npix), 256, 0, static_cast<cudaStream_t>(stream), delta_pre, ld_delta,
             reinterpret_cast<const float2*>(bias), reinterpret_cast<const float2*>(coords0),
             reinterpret_cast<float2*>(coords1), reinterpret_cast<float2*>(flow), flow_pitched, ld_flow, npix);
}
