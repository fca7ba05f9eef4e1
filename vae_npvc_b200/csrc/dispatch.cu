// C-ABI entry points of the three GEMM-shaped ops; picks the tcgen05 or the SIMT backend.
#include "common.cuh"

namespace vnb {
int conv_fwd_simt(const void* in, int P, int B, int Cin, int T, const void* w, int taps, int dil, int Cout,
                  const float* bias, int bias_bstride, const float* residual, float* out, int accumulate,
                  double* gn_sums, int gn_groups, cudaStream_t st);
int conv_wgrad_simt(const void* dy, const void* u, int P, int B, int Cout, int Cin, int T, int taps, int dil,
                    int transposed, float* dw, cudaStream_t st);
int vq_lookup_simt(const float* zn, const float* z_sq, const float* e_hat, const float* e_sq, int B, int D, int T,
                   int K, int64_t* idx, cudaStream_t st);
// tensor-core backend: return VNB_OK, a negative error, or +1 = "shape not taken by this backend"
int conv_fwd_tc(const void* in, int P, int B, int Cin, int T, const void* w, int taps, int dil, int Cout,
                const float* bias, int bias_bstride, const float* residual, float* out, int accumulate,
                double* gn_sums, int gn_groups, cudaStream_t st, const float* mask = nullptr, int mask_act = 0,
                float mask_scale = 1.f);
int conv_wgrad_tc(const void* dy, const void* u, int P, int B, int Cout, int Cin, int T, int taps, int dil,
                  int transposed, float* dw, cudaStream_t st);
int vq_lookup_tc(const void* z_planes, const float* z_sq, const void* e_planes, const float* e_sq, int B, int D,
                 int T, int K, int64_t* idx, cudaStream_t st);
int vq_fused_tc(const float* z, int B, int D, int T, int normalize, const float* e_hat, const void* e_planes, const float* e_sq,
                int K, float* zn_out, float* z_norm_out, int64_t* idx, float* q_out, float* sums, int* counts, cudaStream_t st);
}  // namespace vnb

using namespace vnb;

extern "C" int vnb_conv_fwd(const void* in_planes, int P, int B, int Cin, int T, const void* w_planes, int taps,
                            int dil, int Cout, const float* bias, int bias_bstride, const float* residual, float* out,
                            int accumulate, double* gn_sums, int gn_groups, void* stream) {
    VNB_REQUIRE(in_planes && w_planes && out, "conv_fwd: null pointer");
    VNB_REQUIRE(B > 0 && Cin > 0 && Cout > 0 && T > 0, "conv_fwd: bad shape B=%d Cin=%d Cout=%d T=%d", B, Cin, Cout, T);
    VNB_REQUIRE((P >= 1 && P <= 3) && taps >= 1 && (taps & 1) && dil >= 1, "conv_fwd: P=%d taps=%d dil=%d unsupported", P, taps, dil);
    VNB_REQUIRE(!gn_sums || (gn_groups > 0 && Cout % gn_groups == 0), "conv_fwd: gn_groups=%d does not divide Cout=%d", gn_groups, Cout);
    if (!gn_sums) gn_groups = 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int be = backend();
    if (be != VNB_BACKEND_SIMT) {
        int r = conv_fwd_tc(in_planes, P, B, Cin, T, w_planes, taps, dil, Cout, bias, bias_bstride, residual, out,
                            accumulate, gn_sums, gn_groups, st);
        if (r <= 0) return r;
        if (be == VNB_BACKEND_TC) {
            set_error("conv_fwd: shape not supported by the tcgen05 backend (Cin=%d Cout=%d T=%d)", Cin, Cout, T);
            return VNB_E_ARG;
        }
    }
    return conv_fwd_simt(in_planes, P, B, Cin, T, w_planes, taps, dil, Cout, bias, bias_bstride, residual, out,
                         accumulate, gn_sums, gn_groups, st);
}

// conv + activation backward in the GEMM epilogue: out = mask_scale * conv(in) * act'(mask * mask_scale) + residual.
// The data gradient of a conv whose input went through an activation (EncBlockFn / ConvFn backward) then needs no
// separate vnb_act_bwd_add pass over the result.
extern "C" int vnb_act_bwd_add(const float* du, const float* c, const float* add, int64_t n, int act, float scale,
                               float* out, void* stream);
extern "C" int vnb_conv_fwd_mask(const void* in_planes, int P, int B, int Cin, int T, const void* w_planes, int taps,
                                 int dil, int Cout, const float* residual, float* out, const float* mask, int mask_act,
                                 float mask_scale, void* stream) {
    VNB_REQUIRE(in_planes && w_planes && out && mask, "conv_fwd_mask: null pointer");
    VNB_REQUIRE(B > 0 && Cin > 0 && Cout > 0 && T > 0, "conv_fwd_mask: bad shape B=%d Cin=%d Cout=%d T=%d", B, Cin, Cout, T);
    VNB_REQUIRE((P >= 1 && P <= 3) && taps >= 1 && (taps & 1) && dil >= 1, "conv_fwd_mask: P=%d taps=%d dil=%d unsupported", P, taps, dil);
    VNB_REQUIRE(out != residual && out != mask, "conv_fwd_mask: out must not alias residual / mask");
    cudaStream_t st = (cudaStream_t)stream;
    const int be = backend();
    if (be != VNB_BACKEND_SIMT) {
        int r = conv_fwd_tc(in_planes, P, B, Cin, T, w_planes, taps, dil, Cout, nullptr, 0, residual, out, 0, nullptr, 0, st,
                            mask, mask_act, mask_scale);
        if (r <= 0) return r;
        if (be == VNB_BACKEND_TC) {
            set_error("conv_fwd_mask: shape not supported by the tcgen05 backend (Cin=%d Cout=%d T=%d)", Cin, Cout, T);
            return VNB_E_ARG;
        }
    }
    // SIMT backend: the conv, then the activation backward as its own (in-place, elementwise) pass
    int r = conv_fwd_simt(in_planes, P, B, Cin, T, w_planes, taps, dil, Cout, nullptr, 0, nullptr, out, 0, nullptr, 0, st);
    if (r != VNB_OK) return r;
    return vnb_act_bwd_add(out, mask, residual, (int64_t)B * T * Cout, mask_act, mask_scale, out, stream);
}

extern "C" int vnb_conv_wgrad(const void* dy_planes, const void* u_planes, int P, int B, int Cout, int Cin, int T,
                              int taps, int dil, int transposed, float* dw, void* stream) {
    VNB_REQUIRE(dy_planes && u_planes && dw, "conv_wgrad: null pointer");
    VNB_REQUIRE(B > 0 && Cin > 0 && Cout > 0 && T > 0, "conv_wgrad: bad shape");
    VNB_REQUIRE((P >= 1 && P <= 3) && taps >= 1 && (taps & 1) && dil >= 1, "conv_wgrad: P=%d taps=%d dil=%d unsupported", P, taps, dil);
    cudaStream_t st = (cudaStream_t)stream;
    const int be = backend();
    if (be != VNB_BACKEND_SIMT) {
        int r = conv_wgrad_tc(dy_planes, u_planes, P, B, Cout, Cin, T, taps, dil, transposed, dw, st);
        if (r <= 0) return r;
        if (be == VNB_BACKEND_TC) {
            set_error("conv_wgrad: shape not supported by the tcgen05 backend");
            return VNB_E_ARG;
        }
    }
    return conv_wgrad_simt(dy_planes, u_planes, P, B, Cout, Cin, T, taps, dil, transposed, dw, st);
}

extern "C" int vnb_vq_lookup(const float* zn, const void* z_planes, const float* z_sq, const float* e_hat,
                             const void* e_planes, const float* e_sq, int B, int D, int T, int K, int64_t* idx,
                             void* stream) {
    VNB_REQUIRE(zn && z_sq && e_hat && e_sq && idx, "vq_lookup: null pointer");
    VNB_REQUIRE(B > 0 && D > 0 && T > 0 && K > 0 && D <= 1024, "vq_lookup: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    const int be = backend();
    if (be != VNB_BACKEND_SIMT && z_planes && e_planes) {
        int r = vq_lookup_tc(z_planes, z_sq, e_planes, e_sq, B, D, T, K, idx, st);
        if (r <= 0) return r;
        if (be == VNB_BACKEND_TC) {
            set_error("vq_lookup: shape not supported by the tcgen05 backend (D=%d K=%d)", D, K);
            return VNB_E_ARG;
        }
    }
    return vq_lookup_simt(zn, z_sq, e_hat, e_sq, B, D, T, K, idx, st);
}

// Fused frame preparation + lookup + gather (csrc/tc_vq_fused.cuh).  Returns 1 (no error string) when the shape is not
// taken by the fused kernel (D not in {16, 32, 64, 128}, K < 8, SIMT backend forced): the caller then runs
// vnb_vq_prep_z + vnb_vq_lookup + vnb_vq_gather.
extern "C" int vnb_vq_fused(const float* z, int B, int D, int T, int normalize, const float* e_hat, const void* e_planes,
                            const float* e_sq, int K, float* zn_out, float* z_norm_out, int64_t* idx, float* q_out,
                            float* sums, int* counts, void* stream) {
    VNB_REQUIRE(z && e_hat && e_planes && e_sq && idx, "vq_fused: null pointer");
    VNB_REQUIRE(B > 0 && D > 0 && T > 0 && K > 0, "vq_fused: bad shape");
    VNB_REQUIRE(!normalize || (zn_out && z_norm_out), "vq_fused: the normalised quantiser needs zn_out and z_norm_out");
    if (backend() == VNB_BACKEND_SIMT || option(VNB_OPT_VQ_FUSED) == 0) return 1;
    return vq_fused_tc(z, B, D, T, normalize, e_hat, e_planes, e_sq, K, zn_out, z_norm_out, idx, q_out, sums, counts,
                       (cudaStream_t)stream);
}
