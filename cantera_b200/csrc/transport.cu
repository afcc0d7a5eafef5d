// transport.cu — mixture-averaged transport properties for N states (MixTransport.cpp:33-50,175-202,
// GasTransport.cpp:28-58,60-121,123-146,170-201,288-294): one warp per state, species lane-cyclic.
//   viscosity            Wilke mixture rule over the K x K matrix phi (Poling et al. eq. 9-5.14)
//   thermal conductivity 0.5 (sum x lambda + 1 / sum x / lambda)
//   mixture diffusion    D_k = (W - x_k W_k) / (p W sum_{j != k} x_j / D_jk)
// Species viscosities, conductivities and binary diffusion coefficients come from the reference's fitted
// polynomials in ln T (5 coefficients, non-CK mode). Everything that depends on the molecular weights only
// is tabulated on the host once; the pair loops carry no division except 1 / D_jk.
#include "transport.h"

#include <algorithm>
#include <cmath>
#include <vector>

namespace ctb {

namespace {

constexpr double TR_TINY = 1e-20; // ct_defs.h Tiny: the offset MixTransport::update_C adds to the mole fractions

struct TransportTables {
    int K;
    const double* rmw;   // [K] 1 / W
    const double* mw;    // [K]
    const double* vc;    // [K][5]
    const double* cc;    // [K][5]
    const double* dc;    // [npair][5], pairs i <= j row by row
    const double* wq;    // [hi][lo] (W_lo / W_hi)^(1/4)                  (m_wratjk(k,j), k >= j)
    const double* rden;  // [hi][lo] 1 / (sqrt(8) sqrt(1 + W_hi / W_lo))   (m_wratkj1(j,k), j <= k)
    const double* mwhl;  // [hi][lo] W_hi / W_lo
};

__device__ __forceinline__ double warpSum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ double dot5(const double* p, const double* c)
{
    return p[0] * c[0] + p[1] * c[1] + p[2] * c[2] + p[3] * c[3] + p[4] * c[4]; // utilities.h:63-73
}

template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) mixTransportKernel(const TransportTables t, const TransportArgs a)
{
    extern __shared__ double sm[];
    const int K = t.K;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double* x = sm + (size_t) warp * 5 * K; // mole fractions
    double* sqv = x + K;                    // sqrt(viscosity_k)
    double* rsq = sqv + K;                  // 1 / sqrt(viscosity_k)
    double* cond = rsq + K;
    double* xr = cond + K;                  // scratch: ym
    for (long long n = (long long) blockIdx.x * WARPS + warp; n < a.N; n += (long long) gridDim.x * WARPS) {
        const double T = a.T[n], P = a.P[n];
        const double* Yn = a.Y + n * a.ldY;
        // ---- composition (Phase::setMassFractions, Phase.cpp:342-360; MixTransport::update_C)
        double s = 0.0;
        for (int k = lane; k < K; k += 32) s += fmax(Yn[k], 0.0);
        const double norm = warpSum(s);
        s = 0.0;
        for (int k = lane; k < K; k += 32) {
            const double ym = (fmax(Yn[k], 0.0) / norm) * t.rmw[k];
            xr[k] = ym;
            s += ym;
        }
        const double mmw = 1.0 / warpSum(s);
        const double logt = log(T), sqt = sqrt(T), t14 = sqrt(sqt);
        const double poly[5] = {1.0, logt, logt * logt, logt * logt * logt, logt * logt * logt * logt};
        double s1 = 0.0, s2 = 0.0;
        for (int k = lane; k < K; k += 32) {
            const double xk = fmax(xr[k] * mmw, TR_TINY);
            x[k] = xk;
            const double sq = t14 * dot5(poly, t.vc + 5 * k); // the fit is for sqrt(visc / sqrt(T))
            sqv[k] = sq;
            rsq[k] = 1.0 / sq;
            const double ck = sqt * dot5(poly, t.cc + 5 * k);
            cond[k] = ck;
            s1 += xk * ck;
            s2 += xk / ck;
        }
        s1 = warpSum(s1);
        s2 = warpSum(s2);
        __syncwarp();
        if (a.cond && lane == 0) a.cond[n] = 0.5 * (s1 + 1.0 / s2);
        // ---- viscosity: sum_k x_k mu_k / sum_j phi(k,j) x_j
        if (a.visc) {
            double acc = 0.0;
            for (int k = lane; k < K; k += 32) {
                const double sk = sqv[k], rk = rsq[k];
                double sum = 0.0;
                for (int j = 0; j < K; j++) {
                    const bool low = j <= k; // (k, j) is the directly computed element phi(hi, lo)
                    const int idx = low ? k * K + j : j * K + k;
                    const double shi = low ? sk : sqv[j], rlo = low ? rsq[j] : rk;
                    const double f1 = 1.0 + (shi * rlo) * t.wq[idx];
                    double phi = f1 * f1 * t.rden[idx];
                    if (!low) {
                        // phi(lo, hi) = phi(hi, lo) / ((mu_hi / mu_lo) (W_lo / W_hi))
                        const double vr = (sk * rsq[j]);
                        phi *= (vr * vr) * t.mwhl[idx];
                    }
                    sum += phi * x[j];
                }
                acc += x[k] * (sk * sk) / sum;
            }
            acc = warpSum(acc);
            if (lane == 0) a.visc[n] = acc;
        }
        // ---- mixture-averaged diffusion coefficients
        if (a.dmix) {
            const double tfac = T * sqt;
            for (int k = lane; k < K; k += 32) {
                double sum2 = 0.0;
                for (int j = 0; j < K; j++) {
                    if (j == k) continue;
                    const int i0 = min(j, k), j0 = max(j, k);
                    const int ic = i0 * K - (i0 * (i0 - 1)) / 2 + (j0 - i0);
                    sum2 += x[j] / (tfac * dot5(poly, t.dc + 5 * ic));
                }
                const int ikk = k * K - (k * (k - 1)) / 2;
                const double dkk = tfac * dot5(poly, t.dc + 5 * ikk);
                a.dmix[n * a.ldD + k] = sum2 <= 0.0 ? dkk / P : (mmw - x[k] * t.mw[k]) / (P * mmw * sum2);
            }
        }
        __syncwarp();
    }
}

} // namespace

BatchedMixTransport::BatchedMixTransport(int K, const double* mw, const double* viscPoly, const double* condPoly,
                                         const double* diffPoly, int device)
    : m_K(K), m_device(device)
{
    if (K < 1) throw CtbError("BatchedMixTransport", "no species");
    if (!mw || !viscPoly || !condPoly || !diffPoly) throw CtbError("BatchedMixTransport", "null table pointer");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        throw CtbError("BatchedMixTransport", "no usable CUDA device; this engine has no CPU path");
    }
    if (device < 0 || device >= ndev) throw CtbError("BatchedMixTransport", "device index out of range");
    cudaCheck(cudaSetDevice(device), "cudaSetDevice");
    cudaCheck(cudaDeviceGetAttribute(&m_numSMs, cudaDevAttrMultiProcessorCount, device), "cudaDeviceGetAttribute");
    const size_t npair = (size_t) K * (K + 1) / 2;
    std::vector<double> h;
    h.reserve(2 * K + 10 * K + 5 * npair + 3 * (size_t) K * K);
    for (int k = 0; k < K; k++) h.push_back(1.0 / mw[k]); // Phase::m_rmolwts
    h.insert(h.end(), mw, mw + K);
    h.insert(h.end(), viscPoly, viscPoly + 5 * K);
    h.insert(h.end(), condPoly, condPoly + 5 * K);
    h.insert(h.end(), diffPoly, diffPoly + 5 * npair);
    std::vector<double> wq((size_t) K * K, 0.0), rden((size_t) K * K, 0.0), mwhl((size_t) K * K, 0.0);
    for (int lo = 0; lo < K; lo++) {
        for (int hi = lo; hi < K; hi++) { // GasTransport.cpp:288-294 with j = lo, k = hi
            wq[(size_t) hi * K + lo] = std::sqrt(std::sqrt(mw[lo] / mw[hi]));
            rden[(size_t) hi * K + lo] = 1.0 / (std::sqrt(8.0) * std::sqrt(1.0 + mw[hi] / mw[lo]));
            mwhl[(size_t) hi * K + lo] = mw[hi] / mw[lo];
        }
    }
    h.insert(h.end(), wq.begin(), wq.end());
    h.insert(h.end(), rden.begin(), rden.end());
    h.insert(h.end(), mwhl.begin(), mwhl.end());
    cudaCheck(cudaMalloc((void**) &m_tab, h.size() * sizeof(double)), "cudaMalloc(transport tables)");
    cudaCheck(cudaMemcpy(m_tab, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice), "cudaMemcpy(transport tables)");
    cudaCheck(cudaStreamCreateWithFlags(&m_stream, cudaStreamNonBlocking), "cudaStreamCreate");
}

BatchedMixTransport::~BatchedMixTransport()
{
    cudaSetDevice(m_device);
    if (m_tab) cudaFree(m_tab);
    if (m_stream) cudaStreamDestroy(m_stream);
}

void BatchedMixTransport::synchronize()
{
    cudaCheck(cudaSetDevice(m_device), "cudaSetDevice");
    cudaCheck(cudaStreamSynchronize(stream()), "cudaStreamSynchronize");
}

void BatchedMixTransport::compute(long long N, const double* T, const double* P, const double* Y, long long ldY,
                                  double* visc, double* cond, double* dmix, long long ldD, int memspace)
{
    const char* where = "BatchedMixTransport::compute";
    const int K = m_K;
    if (N < 0) throw CtbError(where, "negative number of states");
    if (N == 0) return;
    if (!T || !P || !Y) throw CtbError(where, "null state pointer");
    if (ldY < K) throw CtbError(where, "ldY is smaller than the number of species");
    if (!visc && !cond && !dmix) throw CtbError(where, "no output requested");
    if (dmix && ldD < K) throw CtbError(where, "ldD is smaller than the number of species");
    cudaCheck(cudaSetDevice(m_device), "cudaSetDevice");
    const size_t npair = (size_t) K * (K + 1) / 2;
    TransportTables t;
    t.K = K;
    t.rmw = m_tab; t.mw = t.rmw + K; t.vc = t.mw + K; t.cc = t.vc + 5 * K; t.dc = t.cc + 5 * K;
    t.wq = t.dc + 5 * npair; t.rden = t.wq + (size_t) K * K; t.mwhl = t.rden + (size_t) K * K;
    constexpr int WARPS = 8;
    const size_t smem = (size_t) WARPS * 5 * K * sizeof(double);
    if (smem > 200 * 1024) throw CtbError(where, "too many species for the transport kernel");
    auto kern = mixTransportKernel<WARPS>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    auto launch = [&](const TransportArgs& a, cudaStream_t st) {
        const int grid = (int) std::max<long long>(1, std::min<long long>((a.N + WARPS - 1) / WARPS, (long long) m_numSMs * 8));
        kern<<<grid, WARPS * 32, smem, st>>>(t, a);
        cudaCheck(cudaGetLastError(), "mixTransportKernel launch");
    };
    if (memspace == CTB_DEVICE) {
        launch(TransportArgs{N, T, P, Y, ldY, visc, cond, dmix, ldD}, stream());
        return;
    }
    if (memspace != CTB_HOST) throw CtbError(where, "memspace must be CTB_HOST or CTB_DEVICE");
    // host buffers: synchronous chunks through one staging buffer (this is not a tuned path)
    const long long chunk = std::min<long long>(N, 1 << 18);
    const size_t per = 2 + (size_t) K + 2 + (size_t) K;
    m_stage.reserve((size_t) chunk * per * sizeof(double));
    cudaStream_t st = stream();
    for (long long n0 = 0; n0 < N; n0 += chunk) {
        const long long n = std::min(chunk, N - n0);
        double* d = m_stage.as<double>();
        double* dT = d; double* dP = dT + chunk; double* dY = dP + chunk;
        double* dV = dY + chunk * K; double* dC = dV + chunk; double* dD = dC + chunk;
        cudaCheck(cudaMemcpyAsync(dT, T + n0, sizeof(double) * n, cudaMemcpyHostToDevice, st), "H2D T");
        cudaCheck(cudaMemcpyAsync(dP, P + n0, sizeof(double) * n, cudaMemcpyHostToDevice, st), "H2D P");
        cudaCheck(cudaMemcpy2DAsync(dY, sizeof(double) * K, Y + n0 * ldY, sizeof(double) * ldY, sizeof(double) * K, n,
                                    cudaMemcpyHostToDevice, st), "H2D Y");
        launch(TransportArgs{n, dT, dP, dY, K, visc ? dV : nullptr, cond ? dC : nullptr, dmix ? dD : nullptr, K}, st);
        if (visc) cudaCheck(cudaMemcpyAsync(visc + n0, dV, sizeof(double) * n, cudaMemcpyDeviceToHost, st), "D2H");
        if (cond) cudaCheck(cudaMemcpyAsync(cond + n0, dC, sizeof(double) * n, cudaMemcpyDeviceToHost, st), "D2H");
        if (dmix) cudaCheck(cudaMemcpy2DAsync(dmix + n0 * ldD, sizeof(double) * ldD, dD, sizeof(double) * K, sizeof(double) * K, n,
                                              cudaMemcpyDeviceToHost, st), "D2H");
        cudaCheck(cudaStreamSynchronize(st), "cudaStreamSynchronize");
    }
}

} // namespace ctb
