// sgemm.cu -- fp32 GEMM on CUDA cores for the fp32-weight corners of the path:
//   * patch-embed convolutions: conv2d_fp32_VALID / conv3d_fp32_VALID (compute/Convolution.cpp:35-90, :179-240) are an
//     im2col followed by one vec_dot_fp32 per output; here the receptive-field gather happens while the A tile is loaded,
//     so the image is read once and no im2col buffer exists;
//   * the eager attention spelling's Q.K^T and P.V (CPUMatmul::execute -> mat_mul fp32, op/CPUMatmul.cpp:63-83) and
//     Linears whose weights the quantizer keeps in fp32.
// Weights of these ops are kept F32 by the quantizer (tools/quantizer/QuantWriter.cpp:10-35), so the arithmetic type is
// fp32 and stays fp32: C = A.B^T (+bias) with fp32 FMA accumulation; only the summation order differs from the
// reference's 8-lane vec_dot_fp32 (tolerance stated in the tests).  Tile 64x64x16, 256 threads, 4x4 per thread.
#include "common.cuh"

struct AView { // how element (m, k) of A is found
    const float *p;
    long long ld;             // plain: p[m*ld + k]
    int C, T, H, W;           // im2col: input [C][T][H][W] (T = 1 for 2-D)
    int kt, kh, kw, st, sh, sw, ot, oh, ow;
};

template <int MODE> // 0 plain, 1 im2col (2-D and 3-D)
static __device__ __forceinline__ float a_at(const AView &a, int m, int k) {
    if (MODE == 0) return a.p[(size_t)m * a.ld + k];
    const int q = k % a.kw;
    const int p = (k / a.kw) % a.kh;
    const int r = (k / (a.kw * a.kh)) % a.kt;
    const int c = k / (a.kw * a.kh * a.kt);
    const int j = m % a.ow;
    const int i = (m / a.ow) % a.oh;
    const int t = m / (a.ow * a.oh);
    return a.p[(((size_t)c * a.T + t * a.st + r) * a.H + i * a.sh + p) * a.W + j * a.sw + q];
}

// C[m*ldcm + n*ldcn] = sum_k A(m,k) * B[n*ldbn + k*ldbk] + bias[n]; blockIdx.z = batch (strides in elements)
template <int MODE>
__global__ void __launch_bounds__(256) k_sgemm(AView a, long long strideA, const float *__restrict__ Bm, long long ldbn, long long ldbk,
                                               long long strideB, const float *__restrict__ bias, float *__restrict__ Cm, long long ldcm,
                                               long long ldcn, long long strideC, int M, int N, int K) {
    constexpr int BM = 64, BN = 64, BK = 16;
    __shared__ float sA[BK][BM + 4];
    __shared__ float sB[BK][BN + 4];
    a.p += (size_t)blockIdx.z * strideA;
    Bm += (size_t)blockIdx.z * strideB;
    Cm += (size_t)blockIdx.z * strideC;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
    for (int k0 = 0; k0 < K; k0 += BK) {
        // A tile: 64 x 16; thread -> (row = t/4, 4 consecutive k) so plain rows are read 16 B at a time
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int r = t >> 2, kk = (t & 3) * 4 + e;
            const int m = m0 + r, k = k0 + kk;
            sA[kk][r] = (m < M && k < K) ? a_at<MODE>(a, m, k) : 0.f;
        }
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int r = t >> 2, kk = (t & 3) * 4 + e;
            const int n = n0 + r, k = k0 + kk;
            sB[kk][r] = (n < N && k < K) ? Bm[(size_t)n * ldbn + (size_t)k * ldbk] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; kk++) {
            const float4 av = *(const float4 *)&sA[kk][ty * 4];
            const float4 bv = *(const float4 *)&sB[kk][tx * 4];
            const float aa[4] = {av.x, av.y, av.z, av.w}, bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = fmaf(aa[i], bb[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int m = m0 + ty * 4 + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int n = n0 + tx * 4 + j;
            if (n >= N) continue;
            float v = acc[i][j];
            if (bias) v = __fadd_rn(v, bias[n]);
            Cm[(size_t)m * ldcm + (size_t)n * ldcn] = v;
        }
    }
}

MLLMCU_API int mllmcu_sgemm(const float *A, long long lda, long long strideA, const float *B, long long ldbn, long long ldbk,
                            long long strideB, const float *bias, float *C, long long ldc, long long strideC, int M, int N, int K,
                            int batch, void *stream) {
    if (M <= 0 || N <= 0 || K <= 0 || batch <= 0 || batch > 65535) return -1;
    AView a{};
    a.p = A;
    a.ld = lda;
    const dim3 grid((N + 63) / 64, (M + 63) / 64, batch);
    k_sgemm<0><<<grid, 256, 0, (cudaStream_t)stream>>>(a, strideA, B, ldbn, ldbk, strideB, bias, C, ldc, 1, strideC, M, N, K);
    CU_LAUNCH_CHECK();
    return 0;
}

static int conv_launch(const float *x, const float *w, const float *bias, float *y, int B, int C, int T, int H, int W, int O, int kt,
                       int kh, int kw, int st, int sh, int sw, int channels_last, cudaStream_t stream) {
    if (B <= 0 || C <= 0 || O <= 0 || kt <= 0 || kh <= 0 || kw <= 0 || st <= 0 || sh <= 0 || sw <= 0) return -1;
    if (T < kt || H < kh || W < kw || B > 65535) return -1;
    AView a{};
    a.p = x;
    a.C = C; a.T = T; a.H = H; a.W = W;
    a.kt = kt; a.kh = kh; a.kw = kw; a.st = st; a.sh = sh; a.sw = sw;
    a.ot = (T - kt) / st + 1; a.oh = (H - kh) / sh + 1; a.ow = (W - kw) / sw + 1;
    const int M = a.ot * a.oh * a.ow, K = C * kt * kh * kw;
    if (kt == T && kh == H && kw == W) {
        // the kernel covers the whole input (Qwen2-VL's patch embed: every "batch" entry is ONE 3 x 2 x 14 x 14 patch): one output position
        // per entry, the receptive field is the entry itself in memory order -- a plain [B x K] . [O x K]^T product with the entries as rows
        // (one launch of 64-row tiles instead of B launches' worth of tiles with one live row each; same k order, same sums)
        AView pa{};
        pa.p = x;
        pa.ld = K;
        const dim3 g1((O + 63) / 64, (B + 63) / 64, 1);
        k_sgemm<0><<<g1, 256, 0, stream>>>(pa, 0, w, K, 1, 0, bias, y, O, 1, 0, B, O, K);
        CU_LAUNCH_CHECK();
        return 0;
    }
    const dim3 grid((O + 63) / 64, (M + 63) / 64, B);
    // output element (m = position, n = channel): channels-first [B][O][M] -> ldcm 1, ldcn M; channels-last [B][M][O] -> ldcm O, ldcn 1
    k_sgemm<1><<<grid, 256, 0, stream>>>(a, (long long)C * T * H * W, w, K, 1, 0, bias, y, channels_last ? O : 1, channels_last ? 1 : M,
                                         (long long)O * M, M, O, K);
    CU_LAUNCH_CHECK();
    return 0;
}

// conv2d_fp32_VALID: x [B][C][H][W], w [O][C][kh][kw], y [B][O][oh][ow] (or channels_last [B][oh*ow][O], the
// flatten + transpose the ViT graphs apply next, models/vit/modeling_vit.hpp)
MLLMCU_API int mllmcu_conv2d_valid(const float *x, const float *w, const float *bias, float *y, int B, int C, int H, int W, int O, int kh,
                                   int kw, int sh, int sw, int channels_last, void *stream) {
    return conv_launch(x, w, bias, y, B, C, 1, H, W, O, 1, kh, kw, 1, sh, sw, channels_last, (cudaStream_t)stream);
}
// conv3d_fp32_VALID: x [B][C][T][H][W], w [O][C][kt][kh][kw], y BCTHW or (channels_last) BTHWC
MLLMCU_API int mllmcu_conv3d_valid(const float *x, const float *w, const float *bias, float *y, int B, int C, int T, int H, int W, int O,
                                   int kt, int kh, int kw, int st, int sh, int sw, int channels_last, void *stream) {
    return conv_launch(x, w, bias, y, B, C, T, H, W, O, kt, kh, kw, st, sh, sw, channels_last, (cudaStream_t)stream);
}
