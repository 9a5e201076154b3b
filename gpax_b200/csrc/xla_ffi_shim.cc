// XLA FFI handlers (jax.ffi custom calls) over the C ABI of include/gpax_b200.h.
//
// north_star: "a thin C-ABI registered as jax.ffi custom calls".  JAX is NOT installed in this
// image or on the GPU box, and the XLA FFI headers (xla/ffi/api/ffi.h) ship only
// with jaxlib, so this file is COMPILE-GATED: gpax_b200/csrc/build.sh builds
// _lib/libgpax_b200_xla.so from it only when `python -c "import jax; print(jax.ffi.include_dir())"`
// succeeds.  Nothing else in the repository depends on it; the handlers add no arithmetic,
// they only unpack XLA buffers into the plain-pointer calls of libgpax_b200.so.
//
// Reference-side registration:
//   lib = ctypes.CDLL("libgpax_b200_xla.so")
//   for name in ("GpxGram", "GpxPotrf", "GpxTrtri", "GpxMll", "GpxMllGrad", "GpxPredict", "GpxAcqReduce"):
//       jax.ffi.register_ffi_target(name, jax.ffi.pycapsule(getattr(lib, name)), platform="CUDA")
// and the call sites they replace: gpax/kernels.py:93-94 (GpxGram), gpax/gp.py:296 (GpxPotrf),
// gpax/gp.py:297-300 (GpxTrtri + GpxPredict), gpax/gp.py:147-151 (GpxMll; GpxMllGrad is its
// jax.custom_vjp backward), gpax/acquisition.py:260-298 (GpxAcqReduce).
//
// XLA owns every buffer (inputs, results and the scratch results declared by the Python side
// with the sizes of the gpx_*_bytes queries); handlers only enqueue on the stream XLA passes.
#include <cstdint>
#include <cuda_runtime_api.h>

#include "xla/ffi/api/ffi.h"
#include "../../include/gpax_b200.h"

namespace ffi = xla::ffi;
using F64 = ffi::Buffer<ffi::F64>;
using RF64 = ffi::ResultBuffer<ffi::F64>;
using RS32 = ffi::ResultBuffer<ffi::S32>;
using RS64 = ffi::ResultBuffer<ffi::S64>;

static ffi::Error status(int rc) {
    if (rc == 0) return ffi::Error::Success();
    return ffi::Error(rc < 0 ? ffi::ErrorCode::kInvalidArgument : ffi::ErrorCode::kInternal,
                      gpx_last_error_string());
}

// kernel description from the per-sample arrays XLA hands over: length [S, dl], period [S, dp] or [0]
static gpx_kernel_params make_kp(int64_t kind, int64_t d, const F64 &scale, const F64 &length,
                                 const F64 &period) {
    gpx_kernel_params kp;
    kp.kind = (int32_t)kind;
    kp.d = (int32_t)d;
    kp.dl = (int32_t)(length.dimensions().size() > 1 ? length.dimensions()[1] : 1);
    const bool has_p = period.element_count() > 0;
    kp.dp = has_p ? (int32_t)(period.dimensions().size() > 1 ? period.dimensions()[1] : 1) : 0;
    kp.scale = scale.typed_data();
    kp.length = length.typed_data();
    kp.period = has_p ? period.typed_data() : nullptr;
    return kp;
}

// out[S, n, ldo] = k(x1, x2; theta_s) (+ diagonal term)        gpax/kernels.py:93-94
static ffi::Error GramImpl(cudaStream_t stream, F64 x1, F64 x2, F64 scale, F64 length, F64 period,
                           F64 noise, F64 jitter, RF64 out, int64_t kind, int64_t add_diag,
                           int64_t noise_is_vector, int64_t region) {
    const int64_t n = x1.dimensions()[0], m = x2.dimensions()[0], d = x1.dimensions()[1];
    const auto od = out->dimensions();                      // [S, n, ldo]
    gpx_kernel_params kp = make_kp(kind, d, scale, length, period);
    return status(gpx_gram(&kp, od[0], n, m, x1.typed_data(), x2.typed_data(),
                           noise.element_count() ? noise.typed_data() : nullptr,
                           noise_is_vector ? 1 : 0, 0, jitter.typed_data(), (int32_t)add_diag,
                           (int32_t)region, out->typed_data(), od[2], stream));
}

// in-place Cholesky: XLA aliases the operand to the first result (input_output_aliases={0: 0})
static ffi::Error PotrfImpl(cudaStream_t stream, F64 a, RF64 l, RF64 dinv, RS32 info, int64_t n) {
    const auto ad = a.dimensions();                         // [S, n, lda]
    if (l->typed_data() != a.typed_data())
        cudaMemcpyAsync(l->typed_data(), a.typed_data(), a.size_bytes(), cudaMemcpyDeviceToDevice, stream);
    return status(gpx_potrf_batched(ad[0], n, l->typed_data(), ad[2], dinv->typed_data(),
                                    info->typed_data(), stream));
}

static ffi::Error TrtriImpl(cudaStream_t stream, F64 l, F64 dinv, RF64 w, int64_t n) {
    const auto ld = l.dimensions();
    return status(gpx_trtri_batched(ld[0], n, l.typed_data(), ld[2], dinv.typed_data(), w->typed_data(),
                                    w->dimensions()[2], stream));
}

// resid [n] (shared) or [S, n]; ws sized by gpx_mll_ws_bytes            gpax/gp.py:147-151,300
static ffi::Error MllImpl(cudaStream_t stream, F64 l, F64 dinv, F64 resid, RF64 alpha, RF64 mll, RF64 ws,
                          int64_t n) {
    const auto ld = l.dimensions();
    const int64_t stride = resid.dimensions().size() > 1 ? n : 0;
    return status(gpx_mll(ld[0], n, l.typed_data(), ld[2], dinv.typed_data(), resid.typed_data(), stride,
                          alpha->typed_data(), mll->typed_data(), ws->typed_data(), stream));
}

// backward of MllImpl under jax.custom_vjp                                  gpax/gp.py:515-530
static ffi::Error MllGradImpl(cudaStream_t stream, F64 x, F64 w, F64 alpha, F64 scale, F64 length,
                              F64 period, RF64 dscale, RF64 dlength, RF64 dperiod, RF64 dnoise, RF64 ws,
                              int64_t kind) {
    const auto wd = w.dimensions();
    gpx_kernel_params kp = make_kp(kind, x.dimensions()[1], scale, length, period);
    return status(gpx_mll_grad(&kp, wd[0], x.dimensions()[0], x.typed_data(), w.typed_data(), wd[2],
                               alpha.typed_data(), dscale->typed_data(), dlength->typed_data(),
                               kp.dp ? dperiod->typed_data() : nullptr, dnoise->typed_data(),
                               ws->typed_data(), stream));
}

// fused predictive (diag of the posterior covariance)                       gpax/gp.py:277-301
static ffi::Error PredictImpl(cudaStream_t stream, F64 x, F64 xnew, F64 w, F64 alpha, F64 scale,
                              F64 length, F64 period, F64 m_new, F64 diag_add, RF64 mu, RF64 var, RF64 ws,
                              int64_t kind, double m_const) {
    const auto wd = w.dimensions();                         // [S, n, ldw]
    gpx_kernel_params kp = make_kp(kind, x.dimensions()[1], scale, length, period);
    return status(gpx_predict(&kp, wd[0], x.dimensions()[0], xnew.dimensions()[0], x.typed_data(),
                              xnew.typed_data(), w.typed_data(), wd[2], alpha.typed_data(),
                              m_new.element_count() ? m_new.typed_data() : nullptr, m_const,
                              diag_add.typed_data(), mu->typed_data(), var->typed_data(),
                              ws->element_count() ? ws->typed_data() : nullptr, stream));
}

// mixture + EI / UCB / UE + per-block winners, then the final first-max       gpax/acquisition.py:214,260-298
static ffi::Error AcqReduceImpl(cudaStream_t stream, F64 mu, F64 var, RF64 out_mu, RF64 out_sd, RF64 out_acq,
                                RF64 part_val, RS64 part_idx, RF64 best_val, RS64 best_idx, int64_t acq,
                                double beta, double y_max, double mu_scale, double mu_shift,
                                double sd_scale, int64_t m_offset) {
    const auto md = mu.dimensions();                        // [S, M]
    int rc = gpx_acq_reduce(md[0], md[1], mu.typed_data(), var.typed_data(), (int32_t)acq, beta, y_max,
                            mu_scale, mu_shift, sd_scale, m_offset, out_mu->typed_data(),
                            out_sd->typed_data(), out_acq->typed_data(), part_val->typed_data(),
                            part_idx->typed_data(), stream);
    if (rc) return status(rc);
    return status(gpx_argmax_finalize(gpx_acq_parts(md[1]), part_val->typed_data(), part_idx->typed_data(),
                                      best_val->typed_data(), best_idx->typed_data(), stream));
}

#define GPX_STREAM ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()

XLA_FFI_DEFINE_HANDLER_SYMBOL(GpxGram, GramImpl,
    GPX_STREAM.Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>()
        .Ret<F64>().Attr<int64_t>("kind").Attr<int64_t>("add_diag").Attr<int64_t>("noise_is_vector")
        .Attr<int64_t>("region"));
XLA_FFI_DEFINE_HANDLER_SYMBOL(GpxPotrf, PotrfImpl,
    GPX_STREAM.Arg<F64>().Ret<F64>().Ret<F64>().Ret<ffi::Buffer<ffi::S32>>().Attr<int64_t>("n"));
XLA_FFI_DEFINE_HANDLER_SYMBOL(GpxTrtri, TrtriImpl,
    GPX_STREAM.Arg<F64>().Arg<F64>().Ret<F64>().Attr<int64_t>("n"));
XLA_FFI_DEFINE_HANDLER_SYMBOL(GpxMll, MllImpl,
    GPX_STREAM.Arg<F64>().Arg<F64>().Arg<F64>().Ret<F64>().Ret<F64>().Ret<F64>().Attr<int64_t>("n"));
XLA_FFI_DEFINE_HANDLER_SYMBOL(GpxMllGrad, MllGradImpl,
    GPX_STREAM.Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>()
        .Ret<F64>().Ret<F64>().Ret<F64>().Ret<F64>().Ret<F64>().Attr<int64_t>("kind"));
XLA_FFI_DEFINE_HANDLER_SYMBOL(GpxPredict, PredictImpl,
    GPX_STREAM.Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>().Arg<F64>()
        .Arg<F64>().Ret<F64>().Ret<F64>().Ret<F64>().Attr<int64_t>("kind").Attr<double>("m_const"));
XLA_FFI_DEFINE_HANDLER_SYMBOL(GpxAcqReduce, AcqReduceImpl,
    GPX_STREAM.Arg<F64>().Arg<F64>().Ret<F64>().Ret<F64>().Ret<F64>().Ret<F64>()
        .Ret<ffi::Buffer<ffi::S64>>().Ret<F64>().Ret<ffi::Buffer<ffi::S64>>()
        .Attr<int64_t>("acq").Attr<double>("beta").Attr<double>("y_max").Attr<double>("mu_scale")
        .Attr<double>("mu_shift").Attr<double>("sd_scale").Attr<int64_t>("m_offset"));
