// api.cu -- error plumbing, launch accounting and the plan runner of libede_sm100.so.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace ede {

static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return EDE_OK;
    set_error("%s: %s", what, cudaGetErrorString(e));
    return EDE_E_CUDA;
}

void count_launch(int n) { g_launches += n; }

}  // namespace ede

extern "C" {

int ede_version(void) { return EDE_ABI_VERSION; }
const char* ede_last_error(void) { return ede::g_err; }

int64_t ede_launch_count(int32_t reset) {
    int64_t v = ede::g_launches;
    if (reset) ede::g_launches = 0;
    return v;
}

int ede_fill_zero(void* p, int64_t bytes, void* stream) {
    EDE_REQUIRE(p != nullptr && bytes >= 0, "ede_fill_zero: bad arguments");
    if (bytes == 0) return EDE_OK;
    ede::count_launch();
    return ede::check_cuda(cudaMemsetAsync(p, 0, (size_t)bytes, (cudaStream_t)stream), "ede_fill_zero");
}

int ede_plan_run(const EdeOp* ops, int32_t n_ops, void* stream) {
    EDE_REQUIRE(ops != nullptr || n_ops == 0, "ede_plan_run: null op list");
    for (int i = 0; i < n_ops; ++i) {
        const EdeOp& op = ops[i];
        int rc;
        switch (op.kind) {
            case EDE_OP_CONV: rc = ede_conv2d(&op.u.conv, stream); break;
            case EDE_OP_DWCONV: rc = ede_dwconv2d(&op.u.dw, stream); break;
            case EDE_OP_SE_FC: rc = ede_se_fc(&op.u.se_fc, stream); break;
            case EDE_OP_SE_SCALE: rc = ede_se_scale(&op.u.se_scale, stream); break;
            case EDE_OP_RESIZE: rc = ede_resize_bilinear_ac(&op.u.resize, stream); break;
            case EDE_OP_MAXPOOL: rc = ede_maxpool3x3s2(&op.u.maxpool, stream); break;
            case EDE_OP_ELTWISE: rc = ede_eltwise(&op.u.eltwise, stream); break;
            case EDE_OP_TO_NHWC: rc = ede_nchw_to_nhwc(&op.u.layout, stream); break;
            case EDE_OP_TO_NCHW: rc = ede_nhwc_to_nchw(&op.u.layout, stream); break;
            case EDE_OP_UPCONV: rc = ede_upconv_taps(&op.u.upconv, stream); break;
            case EDE_OP_STEM: rc = ede_stem_conv(&op.u.stem, stream); break;
            case EDE_OP_TAPSUM: rc = ede_tapsum(&op.u.tapsum, stream); break;
            case EDE_OP_ZERO: rc = ede_fill_zero(op.u.zero.p, op.u.zero.bytes, stream); break;
            default:
                ede::set_error("ede_plan_run: op %d has unknown kind %d", i, op.kind);
                return EDE_E_ARG;
        }
        if (rc != EDE_OK) {
            char prev[400];
            strncpy(prev, ede_last_error(), sizeof(prev) - 1);
            prev[sizeof(prev) - 1] = 0;
            ede::set_error("ede_plan_run: op %d (kind %d) failed: %s", i, op.kind, prev);
            return rc;
        }
    }
    return EDE_OK;
}

}  // extern "C"
