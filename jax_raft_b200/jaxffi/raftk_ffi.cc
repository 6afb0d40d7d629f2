// XLA FFI (jax.ffi) handlers over the raftk C ABI — the custom-call layer BASELINE.json's
// north-star asks for.  SOURCE-COMPLETE BUT NEVER COMPILED OR RUN ON THE BUILD IMAGE: jax, jaxlib
// and the xla/ffi/api headers are absent there.  Build, where jax is installed:
//
//   g++ -O2 -std=c++17 -fPIC -shared -I$(python -c "import jax.ffi; print(jax.ffi.include_dir())") \
//       -I<repo>/include -I/usr/local/cuda/include raftk_ffi.cc -o libraftk_ffi.so \
//       -L<repo>/jax_raft_b200/lib -lraftk -Wl,-rpath,<repo>/jax_raft_b200/lib
//
// Every handler is launch-only on XLA's stream, allocates nothing (scratch is an extra result
// buffer that XLA allocates) and maps a non-zero raftk status to an XLA_FFI_Error.

#include <cstdint>

#include "cuda_runtime_api.h"
#include "raftk.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

static ffi::Error Status(int rc) {
  return rc == 0 ? ffi::Error::Success() : ffi::Error(ffi::ErrorCode::kInternal, raftk_error_string(rc));
}

// ---- CorrBlock.build_pyramid (model.py:418-446) -------------------------------------------------
// fmap1 (B, hq, w, D) [hq = h, or a band of query rows], fmap2 (B, h, w, D)
// -> levels l = 0..3: (B*hq*w, h>>l, w>>l, 1);  scratch: u8[raftk_corr_pyramid_scratch_bytes]
static ffi::Error PyramidImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> fmap1, ffi::Buffer<ffi::F32> fmap2,
                              ffi::ResultBuffer<ffi::F32> l0, ffi::ResultBuffer<ffi::F32> l1,
                              ffi::ResultBuffer<ffi::F32> l2, ffi::ResultBuffer<ffi::F32> l3,
                              ffi::ResultBuffer<ffi::U8> scratch) {
  auto d1 = fmap1.dimensions();
  auto d2 = fmap2.dimensions();
  const int B = (int)d2[0], h = (int)d2[1], w = (int)d2[2], D = (int)d2[3];
  const int nq = (int)(d1[1] * d1[2]);
  float* levels[4] = {l0->typed_data(), l1->typed_data(), l2->typed_data(), l3->typed_data()};
  return Status(raftk_corr_pyramid(fmap1.typed_data(), fmap2.typed_data(), levels, 4, B, nq, h, w, D,
                                   scratch->typed_data(), scratch->element_count(), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(RaftkCorrPyramid, PyramidImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::U8>>());

// ---- CorrBlock.index_pyramid (model.py:448-470) -------------------------------------------------
static ffi::Error LookupImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> l0, ffi::Buffer<ffi::F32> l1,
                             ffi::Buffer<ffi::F32> l2, ffi::Buffer<ffi::F32> l3, ffi::Buffer<ffi::F32> coords,
                             ffi::ResultBuffer<ffi::F32> out, int32_t radius) {
  const float* levels[4] = {l0.typed_data(), l1.typed_data(), l2.typed_data(), l3.typed_data()};
  auto d = l0.dimensions();  // (Q, h, w, 1)
  return Status(raftk_corr_lookup(levels, 4, coords.typed_data(), out->typed_data(), (int)d[0], (int)d[1], (int)d[2],
                                  radius, stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(RaftkCorrLookup, LookupImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Attr<int32_t>("radius"));

// ---- upsample_flow, mask branch (model.py:85-98) ------------------------------------------------
static ffi::Error UpsampleImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> flow, ffi::Buffer<ffi::F32> mask,
                               ffi::ResultBuffer<ffi::F32> out) {
  auto d = flow.dimensions();  // (B, h, w, 2)
  return Status(raftk_convex_upsample(flow.typed_data(), mask.typed_data(), out->typed_data(), (int)d[0], (int)d[1],
                                      (int)d[2], stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(RaftkConvexUpsample, UpsampleImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

// ---- grid_sample (model.py:24-34) ---------------------------------------------------------------
static ffi::Error GridSampleImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> img, ffi::Buffer<ffi::F32> grid,
                                 ffi::ResultBuffer<ffi::F32> out) {
  auto di = img.dimensions();   // (n, h, w, c)
  auto dg = grid.dimensions();  // (n, gh, gw, 2)
  return Status(raftk_grid_sample(img.typed_data(), grid.typed_data(), out->typed_data(), (int)di[0], (int)di[1],
                                  (int)di[2], (int)di[3], (int)dg[1], (int)dg[2], stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(RaftkGridSample, GridSampleImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

// ---- alt_corr (no reference counterpart) --------------------------------------------------------
// pooled levels of fmap2: results l1..l3 (B, h>>l, w>>l, D); level 0 is the input itself
static ffi::Error AltPoolImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> fmap2, ffi::ResultBuffer<ffi::F32> l1,
                              ffi::ResultBuffer<ffi::F32> l2, ffi::ResultBuffer<ffi::F32> l3) {
  auto d = fmap2.dimensions();
  float* levels[4] = {const_cast<float*>(fmap2.typed_data()), l1->typed_data(), l2->typed_data(), l3->typed_data()};
  return Status(raftk_altcorr_pool_fmap2(levels, 4, (int)d[0], (int)d[1], (int)d[2], (int)d[3], stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(RaftkAltcorrPool, AltPoolImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

static ffi::Error AltLookupImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> fmap1, ffi::Buffer<ffi::F32> f0,
                                ffi::Buffer<ffi::F32> f1, ffi::Buffer<ffi::F32> f2, ffi::Buffer<ffi::F32> f3,
                                ffi::Buffer<ffi::F32> coords, ffi::ResultBuffer<ffi::F32> out, int32_t radius) {
  auto dq = fmap1.dimensions();  // (B, hq, w, D)
  auto dt = f0.dimensions();     // (B, h, w, D)
  const float* levels[4] = {f0.typed_data(), f1.typed_data(), f2.typed_data(), f3.typed_data()};
  return Status(raftk_altcorr_lookup(fmap1.typed_data(), levels, 4, coords.typed_data(), out->typed_data(), (int)dt[0],
                                     (int)(dq[1] * dq[2]), (int)dt[1], (int)dt[2], (int)dt[3], radius, stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(RaftkAltcorrLookup, AltLookupImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Attr<int32_t>("radius"));
