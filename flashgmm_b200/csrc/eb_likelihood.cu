// eb_likelihood.cu -- EntropyBottleneck._likelihood (hyper-latent z) as ONE kernel (sm_100a).
//
// The reference evaluates the factorised density with ~30 small tensor ops per call
// (/root/reference/compressai/entropy_models/entropy_models.py:421-463: for each of the 5 layers a batched matmul with
// softplus(matrix), a bias add and a tanh gate, twice -- at x - 1/2 and x + 1/2 -- then two sigmoids and a difference).
// z is tiny (192 x 8 x 12 per 768x512 image), so those calls are pure launch latency. Here: one thread per element,
// the channel's 58 transformed parameters (softplus of the matrices, tanh of the factors, biases; filters (3, 3, 3, 3))
// in shared memory, both chains in registers.
//   values : (C, N) float32, channel-major (what EntropyBottleneck.forward builds, :479-481)
//   params : (C, 58): sp0[3] b0[3] tf0[3] | sp1[9] b1[3] tf1[3] | sp2[9] b2[3] tf2[3] | sp3[9] b3[3] tf3[3] | sp4[3] b4[1]
//   out    : (C, N) sigmoid(logits(v + 1/2)) - sigmoid(logits(v - 1/2))     (the likelihood bound is applied by the caller)
// fp32 with tanhf / expf: 1e-5 relative against the reference's tensor ops (tests/test_model_gpu.py).
#include "gmm_kernels.h"

namespace flashgmm {

namespace {

constexpr int kEbParams = 58;

__device__ __forceinline__ float eb_chain(float v, const float* __restrict__ p) {
    float a[3], b[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {                           // layer 0: 1 -> 3
        const float t = fmaf(p[j], v, p[3 + j]);
        a[j] = fmaf(p[6 + j], tanhf(t), t);
    }
    p += 9;
#pragma unroll
    for (int layer = 0; layer < 3; ++layer) {               // layers 1..3: 3 -> 3
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            float t = p[9 + j];
            t = fmaf(p[3 * j + 0], a[0], t);
            t = fmaf(p[3 * j + 1], a[1], t);
            t = fmaf(p[3 * j + 2], a[2], t);
            b[j] = fmaf(p[12 + j], tanhf(t), t);
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) a[j] = b[j];
        p += 15;
    }
    float t = p[3];                                         // layer 4: 3 -> 1, no gate
    t = fmaf(p[0], a[0], t);
    t = fmaf(p[1], a[1], t);
    return fmaf(p[2], a[2], t);
}

__device__ __forceinline__ float eb_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void __launch_bounds__(256)
eb_likelihood_kernel(const float* __restrict__ values, const float* __restrict__ params, int64_t N, float* __restrict__ out) {
    __shared__ float p[kEbParams];
    const int64_t c = blockIdx.y;
    if (threadIdx.x < kEbParams) p[threadIdx.x] = __ldg(params + c * kEbParams + threadIdx.x);
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = __ldg(values + c * N + i);
        const float lower = eb_chain(v - 0.5f, p), upper = eb_chain(v + 0.5f, p);
        out[c * N + i] = eb_sigmoid(upper) - eb_sigmoid(lower);
    }
}

}  // namespace

cudaError_t launch_eb_likelihood(const float* values, const float* params, int64_t C, int64_t N, float* out, cudaStream_t st) {
    if (C <= 0 || N <= 0) return cudaSuccess;
    if (C > 65535) return cudaErrorInvalidValue;
    const unsigned gx = (unsigned)std::min<int64_t>((N + 255) / 256, 256);
    eb_likelihood_kernel<<<dim3(gx, (unsigned)C), 256, 0, st>>>(values, params, N, out);
    return cudaGetLastError();
}

}  // namespace flashgmm
