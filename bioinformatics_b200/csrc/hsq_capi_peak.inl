// hsq_capi_peak.inl -- measurement helpers: stream timer and the INT32 issue-rate micro-kernel that serves as the
// DP kernels' roofline denominator ("measured sustained INT32 rate from an IADD/IMNMX micro-kernel").
namespace hsq {

// 8 INDEPENDENT dependency chains per thread, half on the FMA pipe (IMAD with run-time multiplier and addend: nothing to fold),
// half on the ALU pipe (VIADDMNMX, the DP cell's own instruction; a LOP3 chain over loop-invariant operands would collapse into a
// single LOP3).  Every loop trip issues 64 integer instructions per thread (+ 3 of loop overhead); 64 resident warps per SM hide the 4-cycle latency.
__global__ void __launch_bounds__(256) int32_peak_kernel(int iters, int seed, int* __restrict__ sink) {
    const int k0 = seed | 3, k1 = seed + 5, k2 = seed ^ 9, k3 = seed + 13;  // run-time values: not immediates
    int a0 = seed + threadIdx.x, a1 = a0 ^ 0x55, a2 = a0 + 7, a3 = a0 * 3, b0 = a0 + 11, b1 = a0 ^ 0x33, b2 = a0 - 5, b3 = a0 + 1;
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            a0 = a0 * k0 + k1;
            b0 = __viaddmax_s32(b0, k1, k2);
            a1 = a1 * k1 + k2;
            b1 = __viaddmax_s32(b1, k2, k3);
            a2 = a2 * k2 + k3;
            b2 = __viaddmax_s32(b2, k3, k0);
            a3 = a3 * k3 + k0;
            b3 = __viaddmax_s32(b3, k0, k1);
        }
    }
    int r = a0 ^ a1 ^ a2 ^ a3 ^ b0 ^ b1 ^ b2 ^ b3;
    if (r == 0x7fffffff) sink[0] = r;  // keeps the chains alive
}

}  // namespace hsq

extern "C" int32_t hsq_timer_start(hsq_ctx* ctx) {
    if (!ctx) return HSQ_ERR_INVALID_ARGUMENT;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    HSQ_CUDA(ctx, cudaEventRecord(ctx->ev_user0, ctx->stream));
    return HSQ_OK;
}

extern "C" int32_t hsq_timer_stop(hsq_ctx* ctx, float* ms) {
    if (!ctx || !ms) return HSQ_ERR_INVALID_ARGUMENT;
    HSQ_CUDA(ctx, cudaEventRecord(ctx->ev_user1, ctx->stream));
    HSQ_CUDA(ctx, cudaEventSynchronize(ctx->ev_user1));
    HSQ_CUDA(ctx, cudaEventElapsedTime(ms, ctx->ev_user0, ctx->ev_user1));
    return HSQ_OK;
}

// Sustained INT32 instruction issue rate of this GPU in lane-operations per second (seconds-long loop, so it sees the
// same power-capped clocks as a long DP step).
extern "C" int32_t hsq_int32_peak(hsq_ctx* ctx, double seconds, double* int_ops_per_s) {
    if (!ctx || !int_ops_per_s) return HSQ_ERR_INVALID_ARGUMENT;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    HSQ_CUDA(ctx, ctx->misc.ensure(256));
    const int grid = ctx->sm_count * 8, block = 256, iters = 1 << 14;
    const double ops_per_launch = (double)grid * block * (double)iters * 64.0;
    int32_peak_kernel<<<grid, block, 0, ctx->stream>>>(1024, 1, ctx->misc.as<int>());  // warm-up
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    double total_ms = 0, total_ops = 0, best = 0;
    while (total_ms < seconds * 1e3) {
        HSQ_CUDA(ctx, cudaEventRecord(ctx->ev_user0, ctx->stream));
        int32_peak_kernel<<<grid, block, 0, ctx->stream>>>(iters, 1, ctx->misc.as<int>());
        HSQ_CUDA(ctx, cudaEventRecord(ctx->ev_user1, ctx->stream));
        HSQ_CUDA(ctx, cudaEventSynchronize(ctx->ev_user1));
        float ms = 0;
        HSQ_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_user0, ctx->ev_user1));
        total_ms += ms;
        total_ops += ops_per_launch;
        best = std::max(best, ops_per_launch / (ms * 1e-3));
    }
    (void)best;
    *int_ops_per_s = total_ops / (total_ms * 1e-3);
    return HSQ_OK;
}
