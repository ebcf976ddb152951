// pose.cu — batched pose post-processing on the device: what the reference does on the host, one 3x3 at a
// time in NumPy float64, right after the model when the `se3` output is used:
//   pcp_dcm_from_rpy_f32      <- get_dcm                  (PCA/utils/mat_ops.py:6-39)
//   pcp_rpy_from_dcm_f32      <- get_roll_pitch_yaw_deg   (PCA/utils/mat_ops.py:30-36)
//   pcp_reorthogonalize_f32   <- reorthogonalize          (PCA/utils/mat_ops.py:63-65)   U V^T of the SVD, no determinant fix
//   pcp_kabsch_f32            <- solve_kabsch             (PCT/utils/calibration.py:3-31)
// fp32 in / out (the model's R is fp32), fp64 arithmetic inside like the reference's NumPy.  One thread per matrix; Kabsch
// one CTA per point-set pair.  3x3 SVD: cyclic Jacobi on M^T M for V and the singular values, U from M V with the third
// column completed as a signed cross product (exactly orthonormal also when the smallest singular value is tiny).
#include "common.cuh"

namespace pcp {

struct Svd3 { double U[9], V[9], s[3]; };          // row-major; M = U diag(s) V^T, s descending

__device__ void svd3(const double *M, Svd3 &o)
{
    double S[9], V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) S[i * 3 + j] = M[0 * 3 + i] * M[0 * 3 + j] + M[1 * 3 + i] * M[1 * 3 + j] + M[2 * 3 + i] * M[2 * 3 + j];
    for (int sweep = 0; sweep < 12; ++sweep) {
        const double off = fabs(S[1]) + fabs(S[2]) + fabs(S[5]);
        if (off <= 1e-300 || off <= 1e-17 * (fabs(S[0]) + fabs(S[4]) + fabs(S[8]))) break;
        for (int pq = 0; pq < 3; ++pq) {
            const int p = pq == 2 ? 1 : 0, q = pq == 0 ? 1 : 2;
            const double apq = S[p * 3 + q];
            if (fabs(apq) < 1e-300) continue;
            const double theta = (S[q * 3 + q] - S[p * 3 + p]) / (2.0 * apq);
            const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
            for (int k = 0; k < 3; ++k) {                     // S <- S J (columns p, q)
                const double skp = S[k * 3 + p], skq = S[k * 3 + q];
                S[k * 3 + p] = c * skp - s * skq; S[k * 3 + q] = s * skp + c * skq;
            }
            for (int k = 0; k < 3; ++k) {                     // S <- J^T S (rows p, q)
                const double spk = S[p * 3 + k], sqk = S[q * 3 + k];
                S[p * 3 + k] = c * spk - s * sqk; S[q * 3 + k] = s * spk + c * sqk;
            }
            for (int k = 0; k < 3; ++k) {
                const double vkp = V[k * 3 + p], vkq = V[k * 3 + q];
                V[k * 3 + p] = c * vkp - s * vkq; V[k * 3 + q] = s * vkp + c * vkq;
            }
        }
    }
    int idx[3] = {0, 1, 2};
    double lam[3] = {S[0], S[4], S[8]};
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2 - a; ++b)
            if (lam[idx[b]] < lam[idx[b + 1]]) { const int tmp = idx[b]; idx[b] = idx[b + 1]; idx[b + 1] = tmp; }
    for (int j = 0; j < 3; ++j) {
        o.s[j] = sqrt(fmax(lam[idx[j]], 0.0));
        for (int k = 0; k < 3; ++k) o.V[k * 3 + j] = V[k * 3 + idx[j]];
    }
    double u[3][3];                                           // u[j] = column j of U
    for (int j = 0; j < 3; ++j)
        for (int k = 0; k < 3; ++k) u[j][k] = M[k * 3 + 0] * o.V[0 * 3 + j] + M[k * 3 + 1] * o.V[1 * 3 + j] + M[k * 3 + 2] * o.V[2 * 3 + j];
    auto norm = [](double *v) { const double n = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); if (n > 0) { v[0] /= n; v[1] /= n; v[2] /= n; } return n; };
    norm(u[0]);
    const double d01 = u[0][0] * u[1][0] + u[0][1] * u[1][1] + u[0][2] * u[1][2];
    for (int k = 0; k < 3; ++k) u[1][k] -= d01 * u[0][k];
    norm(u[1]);
    double cr[3] = {u[0][1] * u[1][2] - u[0][2] * u[1][1], u[0][2] * u[1][0] - u[0][0] * u[1][2], u[0][0] * u[1][1] - u[0][1] * u[1][0]};
    const double sgn = (cr[0] * u[2][0] + cr[1] * u[2][1] + cr[2] * u[2][2]) < 0 ? -1.0 : 1.0;
    for (int k = 0; k < 3; ++k) u[2][k] = sgn * cr[k];
    for (int j = 0; j < 3; ++j)
        for (int k = 0; k < 3; ++k) o.U[k * 3 + j] = u[j][k];
}

__device__ __forceinline__ void mat3_abt(const double *A, const double *B, const double *d, double *C)   // C = A diag(d) B^T
{
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) C[i * 3 + j] = A[i * 3 + 0] * d[0] * B[j * 3 + 0] + A[i * 3 + 1] * d[1] * B[j * 3 + 1] + A[i * 3 + 2] * d[2] * B[j * 3 + 2];
}

__global__ void dcm_from_rpy_kernel(const float *__restrict__ rpy, int b, float *__restrict__ R)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= b) return;
    const double k = 3.14159265358979323846 / 180.0;
    const double r = rpy[3 * i] * k, p = rpy[3 * i + 1] * k, y = rpy[3 * i + 2] * k;
    const double cr = cos(r), sr = sin(r), cp = cos(p), sp = sin(p), cy = cos(y), sy = sin(y);
    // A = roll . pitch . yaw (mat_ops.py:6-28), R = A^T
    const double A[9] = {cp * cy, cp * sy, -sp,
                         sr * sp * cy - cr * sy, sr * sp * sy + cr * cy, sr * cp,
                         cr * sp * cy + sr * sy, cr * sp * sy - sr * cy, cr * cp};
    for (int a = 0; a < 3; ++a)
        for (int c = 0; c < 3; ++c) R[9 * i + a * 3 + c] = (float)A[c * 3 + a];
}

__global__ void rpy_from_dcm_kernel(const float *__restrict__ R, int b, float *__restrict__ rpy)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= b) return;
    const double k = 180.0 / 3.14159265358979323846;
    const float *m = R + 9 * i;
    rpy[3 * i] = (float)(atan2((double)m[5], (double)m[8]) * k);          // roll  = atan2(dcm[1][2], dcm[2][2])
    rpy[3 * i + 1] = (float)(-asin((double)m[2]) * k);                    // pitch = -asin(dcm[0][2])
    rpy[3 * i + 2] = (float)(atan2((double)m[1], (double)m[0]) * k);      // yaw   = atan2(dcm[0][1], dcm[0][0])
}

__global__ void reorthogonalize_kernel(const float *__restrict__ R, int b, float *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= b) return;
    double M[9], C[9];
    for (int k = 0; k < 9; ++k) M[k] = (double)R[9 * i + k];
    Svd3 s;
    svd3(M, s);
    const double one[3] = {1, 1, 1};
    mat3_abt(s.U, s.V, one, C);
    for (int k = 0; k < 9; ++k) out[9 * i + k] = (float)C[k];
}

// one CTA per (truth, measured) pair of n x 3 point sets
__global__ void __launch_bounds__(256)
kabsch_kernel(const float *__restrict__ truth, const float *__restrict__ meas, int64_t n, float *__restrict__ Rout, float *__restrict__ tout,
              float *__restrict__ rmse)
{
    __shared__ double red[8][9];
    __shared__ double fin[16];                                  // means (6) ... then R (9) + t (3)
    const int64_t pb = blockIdx.x;
    const float *T = truth + pb * n * 3, *Mm = meas + pb * n * 3;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    auto block_sum = [&](double *v, int cnt, double *dst) {
        for (int k = 0; k < cnt; ++k) {
            double x = v[k];
            for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
            if (lane == 0) red[warp][k] = x;
        }
        __syncthreads();
        if (tid < cnt) { double s = 0; for (int w = 0; w < 8; ++w) s += red[w][tid]; dst[tid] = s; }
        __syncthreads();
    };
    double v[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int64_t i = tid; i < n; i += 256)
        for (int a = 0; a < 3; ++a) { v[a] += (double)T[3 * i + a]; v[3 + a] += (double)Mm[3 * i + a]; }
    block_sum(v, 6, fin);
    double tc[3], mc[3];
    for (int a = 0; a < 3; ++a) { tc[a] = fin[a] / (double)n; mc[a] = fin[3 + a] / (double)n; }
    __syncthreads();
    for (int k = 0; k < 9; ++k) v[k] = 0;
    for (int64_t i = tid; i < n; i += 256)                     // H = meas_centered^T . truth_centered
        for (int a = 0; a < 3; ++a)
            for (int c = 0; c < 3; ++c) v[a * 3 + c] += ((double)Mm[3 * i + a] - mc[a]) * ((double)T[3 * i + c] - tc[c]);
    block_sum(v, 9, fin);
    if (tid == 0) {
        double H[9], R[9];
        for (int k = 0; k < 9; ++k) H[k] = fin[k];
        Svd3 s;
        svd3(H, s);
        double d[3] = {1, 1, 1};
        mat3_abt(s.V, s.U, d, R);                               // R = V U^T
        const double det = R[0] * (R[4] * R[8] - R[5] * R[7]) - R[1] * (R[3] * R[8] - R[5] * R[6]) + R[2] * (R[3] * R[7] - R[4] * R[6]);
        if (det < 0) { d[2] = -1; mat3_abt(s.V, s.U, d, R); }   // Vt[2, :] *= -1 (calibration.py:22-24)
        for (int k = 0; k < 9; ++k) fin[k] = R[k];
        for (int a = 0; a < 3; ++a) fin[9 + a] = tc[a] - (R[a * 3] * mc[0] + R[a * 3 + 1] * mc[1] + R[a * 3 + 2] * mc[2]);
    }
    __syncthreads();
    double R[9], t[3];
    for (int k = 0; k < 9; ++k) R[k] = fin[k];
    for (int a = 0; a < 3; ++a) t[a] = fin[9 + a];
    __syncthreads();
    double e = 0;
    for (int64_t i = tid; i < n; i += 256)
        for (int a = 0; a < 3; ++a) {
            const double r = R[a * 3] * (double)Mm[3 * i] + R[a * 3 + 1] * (double)Mm[3 * i + 1] + R[a * 3 + 2] * (double)Mm[3 * i + 2] + t[a] - (double)T[3 * i + a];
            e += r * r;
        }
    block_sum(&e, 1, fin);
    if (tid == 0) {
        for (int k = 0; k < 9; ++k) Rout[pb * 9 + k] = (float)R[k];
        for (int a = 0; a < 3; ++a) tout[pb * 3 + a] = (float)t[a];
        if (rmse) rmse[pb] = (float)sqrt(fin[0] / (double)n);
    }
}

}  // namespace pcp

using namespace pcp;

extern "C" int pcp_dcm_from_rpy_f32(const float *d_rpy_deg, int b, float *d_R, pcp_stream stream)
{
    PCP_REQUIRE(d_rpy_deg && d_R && b >= 0, PCP_ERR_INVALID, "pcp_dcm_from_rpy_f32: bad arguments");
    if (b == 0) return PCP_OK;
    dcm_from_rpy_kernel<<<(b + 127) / 128, 128, 0, as_stream(stream)>>>(d_rpy_deg, b, d_R);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

extern "C" int pcp_rpy_from_dcm_f32(const float *d_R, int b, float *d_rpy_deg, pcp_stream stream)
{
    PCP_REQUIRE(d_rpy_deg && d_R && b >= 0, PCP_ERR_INVALID, "pcp_rpy_from_dcm_f32: bad arguments");
    if (b == 0) return PCP_OK;
    rpy_from_dcm_kernel<<<(b + 127) / 128, 128, 0, as_stream(stream)>>>(d_R, b, d_rpy_deg);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

extern "C" int pcp_reorthogonalize_f32(const float *d_R, int b, float *d_out, pcp_stream stream)
{
    PCP_REQUIRE(d_R && d_out && b >= 0, PCP_ERR_INVALID, "pcp_reorthogonalize_f32: bad arguments");
    if (b == 0) return PCP_OK;
    reorthogonalize_kernel<<<(b + 63) / 64, 64, 0, as_stream(stream)>>>(d_R, b, d_out);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

extern "C" int pcp_kabsch_f32(const float *d_truth, const float *d_measured, int b, int64_t n, float *d_R, float *d_t, float *d_rmse,
                              pcp_stream stream)
{
    PCP_REQUIRE(d_truth && d_measured && d_R && d_t && b >= 0 && n >= 1, PCP_ERR_INVALID, "pcp_kabsch_f32: bad arguments");
    if (b == 0) return PCP_OK;
    kabsch_kernel<<<b, 256, 0, as_stream(stream)>>>(d_truth, d_measured, n, d_R, d_t, d_rmse);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}
