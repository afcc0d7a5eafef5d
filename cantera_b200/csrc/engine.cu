// engine.cu — BatchedGasKinetics host side: table upload, host<->device staging, launches.
#include "engine.h"
#include "derivs_kernels.h"

#include <algorithm>
#include <cstring>
#include <thread>

namespace ctb {

void cudaCheck(cudaError_t e, const char* where)
{
    if (e != cudaSuccess) {
        throw CtbError(where, std::string("CUDA error: ") + cudaGetErrorString(e));
    }
}

void DevBuf::reserve(size_t n)
{
    if (n <= bytes) return;
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
    cudaCheck(cudaMalloc(&p, n), "DevBuf::reserve");
    bytes = n;
}

DevBuf::~DevBuf()
{
    if (p) cudaFree(p);
}

void PinBuf::reserve(size_t n)
{
    if (n <= bytes) return;
    if (p) cudaFreeHost(p);
    p = nullptr;
    bytes = 0;
    cudaCheck(cudaHostAlloc(&p, n, cudaHostAllocDefault), "PinBuf::reserve");
    bytes = n;
}

PinBuf::~PinBuf()
{
    if (p) cudaFreeHost(p);
}

//! True when the DMA engines can read / write the host range directly (cudaHostAlloc, cudaHostRegister or
//! managed memory); pageable memory reports cudaMemoryTypeUnregistered.
static bool isPinned(const void* ptr)
{
    if (!ptr) return true;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type != cudaMemoryTypeUnregistered;
}

//! rows x width doubles between two host arrays, split over a few threads: one core moves ~10 GB/s of
//! pageable memory, the PCIe link five times that.
static void hostCopyRows(double* dst, long long ldDst, const double* src, long long ldSrc, long long width, long long rows)
{
    auto part = [=](long long r0, long long r1) {
        if (ldDst == width && ldSrc == width) {
            std::memcpy(dst + r0 * width, src + r0 * width, sizeof(double) * width * (r1 - r0));
        } else {
            for (long long r = r0; r < r1; r++) std::memcpy(dst + r * ldDst, src + r * ldSrc, sizeof(double) * width);
        }
    };
    const long long bytes = rows * width * (long long) sizeof(double);
    const int nt = (int) std::min<long long>(4, bytes >> 20);
    if (nt <= 1) {
        part(0, rows);
        return;
    }
    std::thread th[4];
    for (int i = 1; i < nt; i++) th[i] = std::thread(part, rows * i / nt, rows * (i + 1) / nt);
    part(0, rows / nt);
    for (int i = 1; i < nt; i++) th[i].join();
}

BatchedGasKinetics::BatchedGasKinetics(const ctb_mech_tables& tables, int device)
    : m_tables(cantera_b200::MechTables::fromView(tables))
{
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        // There is deliberately no CPU fallback.
        throw CtbError("BatchedGasKinetics", std::string("no usable CUDA device (")
                       + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0")
                       + "); this engine has no CPU path");
    }
    if (device < 0 || device >= ndev) {
        throw CtbError("BatchedGasKinetics", "device index " + std::to_string(device) + " out of range");
    }
    cudaCheck(cudaSetDevice(device), "cudaSetDevice");
    cudaDeviceProp prop;
    cudaCheck(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
    if (prop.major < 10) {
        throw CtbError("BatchedGasKinetics", "built for sm_100a (Blackwell); device is sm_"
                       + std::to_string(prop.major) + std::to_string(prop.minor));
    }
    if (m_tables.n_species < 1) throw CtbError("BatchedGasKinetics", "mechanism has no species");
    try {
        m_packed = packMechanism(m_tables.view());
    } catch (std::exception& ex) {
        throw CtbError("BatchedGasKinetics", ex.what());
    }
    m_dev.view = m_packed.v;
    m_dev.tile = m_packed.tv;
    m_dev.device = device;
    m_dev.numSMs = prop.multiProcessorCount;
    cudaCheck(cudaDeviceGetAttribute(&m_dev.smemOptIn, cudaDevAttrMaxSharedMemoryPerBlockOptin, device),
              "cudaDeviceGetAttribute");
    cudaCheck(cudaMalloc((void**) &m_dev.dBlob, m_packed.blob.size()), "cudaMalloc(mechanism)");
    cudaCheck(cudaMemcpy(m_dev.dBlob, m_packed.blob.data(), m_packed.blob.size(), cudaMemcpyHostToDevice),
              "cudaMemcpy(mechanism)");
    cudaCheck(cudaStreamCreateWithFlags(&m_stream, cudaStreamNonBlocking), "cudaStreamCreate");
    for (int i = 0; i < NSTAGE; i++) {
        cudaCheck(cudaStreamCreateWithFlags(&m_copyStream[i], cudaStreamNonBlocking), "cudaStreamCreate");
        cudaCheck(cudaEventCreateWithFlags(&m_evt[i], cudaEventDisableTiming), "cudaEventCreate");
    }
}

BatchedGasKinetics::~BatchedGasKinetics()
{
    cudaSetDevice(m_dev.device);
    if (m_dev.dBlob) cudaFree(m_dev.dBlob);
    for (int i = 0; i < NSTAGE; i++) {
        if (m_copyStream[i]) cudaStreamDestroy(m_copyStream[i]);
        if (m_evt[i]) cudaEventDestroy(m_evt[i]);
    }
    if (m_stream) cudaStreamDestroy(m_stream);
}

void BatchedGasKinetics::uploadMultipliers()
{
    // the multipliers are folded into several tables (o_mult, the tile region's A and auxMult):
    // repack (same layout, same size) and upload the whole image; setMultiplier is a rare call
    PackedMech again = packMechanism(m_tables.view());
    if (again.blob.size() != m_packed.blob.size()) throw CtbError("BatchedGasKinetics::setMultiplier", "image size changed");
    m_packed = std::move(again);
    m_dev.view = m_packed.v;
    m_dev.tile = m_packed.tv;
    cudaCheck(cudaSetDevice(m_dev.device), "cudaSetDevice");
    cudaCheck(cudaMemcpy(m_dev.dBlob, m_packed.blob.data(), m_packed.blob.size(), cudaMemcpyHostToDevice),
              "cudaMemcpy(multipliers)");
}

void BatchedGasKinetics::setMultiplier(int i, double f)
{
    if (i < 0 || i >= nReactions()) {
        throw CtbError("BatchedGasKinetics::setMultiplier", "reaction index " + std::to_string(i)
                       + " out of range (nReactions = " + std::to_string(nReactions()) + ")");
    }
    synchronize();
    m_tables.multiplier[i] = f;
    uploadMultipliers();
}

double BatchedGasKinetics::multiplier(int i) const
{
    if (i < 0 || i >= nReactions()) {
        throw CtbError("BatchedGasKinetics::multiplier", "reaction index " + std::to_string(i) + " out of range");
    }
    return m_tables.multiplier[i];
}

void BatchedGasKinetics::setDerivativeSettings(bool skipThirdBodies, bool skipFalloff, double rtolDelta)
{
    if (!(rtolDelta > 0.0)) {
        throw CtbError("BatchedGasKinetics::setDerivativeSettings", "rtol-delta must be positive");
    }
    m_skip3b = skipThirdBodies;
    m_skipFalloff = skipFalloff;
    m_rtolDelta = rtolDelta; // kept for API symmetry: the falloff [M]-derivative is analytic here
}

void BatchedGasKinetics::synchronize()
{
    cudaCheck(cudaSetDevice(m_dev.device), "cudaSetDevice");
    cudaCheck(cudaStreamSynchronize(stream()), "cudaStreamSynchronize");
}

// Row-strided copies degrade to one DMA descriptor per row; contiguous batches (ld == width, the
// common case) go through a single linear copy, which runs at PCIe line rate from pinned memory.
static void copyRows(double* dst, long long ldDst, const double* src, long long ldSrc, long long width,
                     long long rows, cudaMemcpyKind kind, cudaStream_t st, const char* what)
{
    if (ldDst == width && ldSrc == width) {
        cudaCheck(cudaMemcpyAsync(dst, src, sizeof(double) * width * rows, kind, st), what);
    } else {
        cudaCheck(cudaMemcpy2DAsync(dst, sizeof(double) * ldDst, src, sizeof(double) * ldSrc,
                                    sizeof(double) * width, rows, kind, st), what);
    }
}

static bool anyFull(const RateOutputs& o)
{
    return o.cdot || o.ddot || o.ropf || o.ropr || o.ropnet || o.kf || o.kc;
}

void BatchedGasKinetics::rates(long long N, const double* T, const double* P, const double* Y,
                               long long ldY, const RateOutputs& out, int memspace)
{
    const int K = nSpecies(), R = nReactions();
    if (N < 0) throw CtbError("BatchedGasKinetics::rates", "negative number of states");
    if (N == 0) return;
    if (!T || !P || !Y) throw CtbError("BatchedGasKinetics::rates", "null state pointer");
    if (ldY < K) throw CtbError("BatchedGasKinetics::rates", "ldY = " + std::to_string(ldY)
                                + " is smaller than the number of species " + std::to_string(K));
    const bool wantW = out.wdot || out.cdot || out.ddot;
    const bool wantR = out.ropf || out.ropr || out.ropnet || out.kf || out.kc;
    if (!wantW && !wantR) throw CtbError("BatchedGasKinetics::rates", "no output requested");
    if (wantW && out.ldW < K) throw CtbError("BatchedGasKinetics::rates", "ldW is smaller than the number of species");
    if (wantR && out.ldR < R) throw CtbError("BatchedGasKinetics::rates", "ldR is smaller than the number of reactions");
    cudaCheck(cudaSetDevice(m_dev.device), "cudaSetDevice");
    const bool full = anyFull(out);

    if (memspace == CTB_DEVICE) {
        RatesArgs a = {N, T, P, Y, ldY, out.wdot, out.cdot, out.ddot, out.ldW,
                       out.ropf, out.ropr, out.ropnet, out.kf, out.kc, out.ldR};
        launchRates(m_dev, a, full, stream());
        cudaCheck(cudaGetLastError(), "ratesKernel launch");
        return;
    }
    if (memspace != CTB_HOST) throw CtbError("BatchedGasKinetics::rates", "memspace must be CTB_HOST or CTB_DEVICE");
    if (!(isPinned(T) && isPinned(P) && isPinned(Y) && isPinned(out.wdot) && isPinned(out.cdot) && isPinned(out.ddot)
          && isPinned(out.ropf) && isPinned(out.ropr) && isPinned(out.ropnet) && isPinned(out.kf) && isPinned(out.kc))) {
        ratesPageable(N, T, P, Y, ldY, out, full);
        return;
    }

    // Pinned host buffers: chunks in flight on NSTAGE streams so that the copies of one chunk overlap the
    // kernel of another.
    double* outsW[3] = {out.wdot, out.cdot, out.ddot};
    double* outsR[5] = {out.ropf, out.ropr, out.ropnet, out.kf, out.kc};
    int nW = 0, nR = 0;
    for (double* p : outsW) nW += p != nullptr;
    for (double* p : outsR) nR += p != nullptr;
    const size_t perState = sizeof(double) * (2 + (size_t) K + (size_t) nW * K + (size_t) nR * R);
    long long chunk = std::max<long long>(1024, std::min<long long>(N, (long long) ((128u << 20) / perState)));
    chunk = std::min<long long>(chunk, std::max<long long>(1024, (N + 31) / 32));
    for (int b = 0; b < NSTAGE; b++) m_stage[b].reserve((size_t) chunk * perState);
    int b = 0;
    for (long long n0 = 0; n0 < N; n0 += chunk, b = (b + 1) % NSTAGE) {
        const long long n = std::min(chunk, N - n0);
        cudaStream_t st = m_copyStream[b];
        double* d = m_stage[b].as<double>();
        double* dT = d; d += chunk;
        double* dP = d; d += chunk;
        double* dY = d; d += chunk * K;
        cudaCheck(cudaMemcpyAsync(dT, T + n0, sizeof(double) * n, cudaMemcpyHostToDevice, st), "H2D T");
        cudaCheck(cudaMemcpyAsync(dP, P + n0, sizeof(double) * n, cudaMemcpyHostToDevice, st), "H2D P");
        copyRows(dY, K, Y + n0 * ldY, ldY, K, n, cudaMemcpyHostToDevice, st, "H2D Y");
        RatesArgs a = {};
        a.N = n; a.T = dT; a.P = dP; a.Y = dY; a.ldY = K; a.ldW = K; a.ldR = R;
        double** slotsW[3] = {&a.wdot, &a.cdot, &a.ddot};
        double** slotsR[5] = {&a.ropf, &a.ropr, &a.ropnet, &a.kf, &a.kc};
        for (int i = 0; i < 3; i++) if (outsW[i]) { *slotsW[i] = d; d += chunk * K; }
        for (int i = 0; i < 5; i++) if (outsR[i]) { *slotsR[i] = d; d += chunk * R; }
        launchRates(m_dev, a, full, st);
        cudaCheck(cudaGetLastError(), "ratesKernel launch");
        for (int i = 0; i < 3; i++) {
            if (outsW[i]) {
                copyRows(outsW[i] + n0 * out.ldW, out.ldW, *slotsW[i], K, K, n, cudaMemcpyDeviceToHost, st, "D2H");
            }
        }
        for (int i = 0; i < 5; i++) {
            if (outsR[i]) {
                copyRows(outsR[i] + n0 * out.ldR, out.ldR, *slotsR[i], R, R, n, cudaMemcpyDeviceToHost, st, "D2H");
            }
        }
    }
    for (int i = 0; i < NSTAGE; i++) cudaCheck(cudaStreamSynchronize(m_copyStream[i]), "cudaStreamSynchronize");
}

// Pageable host buffers: cudaMemcpyAsync from pageable memory is staged by the driver through ONE small
// pinned buffer and blocks the calling thread, so nothing overlaps (15.8 M gri30 states/s). Here every
// stream owns a pinned slot: the host packs chunk i into its slot (threads), the DMA engines and the
// kernel work on it asynchronously, and the results of chunk i - NSTAGE are unpacked when the slot comes
// round again.
void BatchedGasKinetics::ratesPageable(long long N, const double* T, const double* P, const double* Y, long long ldY,
                                       const RateOutputs& out, bool full)
{
    const int K = nSpecies(), R = nReactions();
    double* outsW[3] = {out.wdot, out.cdot, out.ddot};
    double* outsR[5] = {out.ropf, out.ropr, out.ropnet, out.kf, out.kc};
    int nW = 0, nR = 0;
    for (double* p : outsW) nW += p != nullptr;
    for (double* p : outsR) nR += p != nullptr;
    const size_t inDoubles = 2 + (size_t) K, outDoubles = (size_t) nW * K + (size_t) nR * R;
    const size_t perState = sizeof(double) * (inDoubles + outDoubles);
    long long chunk = std::max<long long>(1024, std::min<long long>(N, (long long) ((32u << 20) / perState)));
    chunk = std::min<long long>(chunk, std::max<long long>(1024, (N + 15) / 16));
    for (int b = 0; b < NSTAGE; b++) {
        m_stage[b].reserve((size_t) chunk * perState);
        m_pin[b].reserve((size_t) chunk * perState);
    }
    long long pendN0[NSTAGE], pendN[NSTAGE];
    for (int b = 0; b < NSTAGE; b++) pendN[b] = 0;
    // slot layout (device and pinned alike): T[chunk] P[chunk] Y[chunk][K] | outputs, each [chunk][width]
    auto drain = [&](int b) {
        if (!pendN[b]) return;
        cudaCheck(cudaStreamSynchronize(m_copyStream[b]), "cudaStreamSynchronize");
        const double* h = m_pin[b].as<double>() + (size_t) chunk * inDoubles;
        for (int i = 0; i < 3; i++) {
            if (outsW[i]) { hostCopyRows(outsW[i] + pendN0[b] * out.ldW, out.ldW, h, K, K, pendN[b]); h += chunk * K; }
        }
        for (int i = 0; i < 5; i++) {
            if (outsR[i]) { hostCopyRows(outsR[i] + pendN0[b] * out.ldR, out.ldR, h, R, R, pendN[b]); h += chunk * R; }
        }
        pendN[b] = 0;
    };
    int b = 0;
    for (long long n0 = 0; n0 < N; n0 += chunk, b = (b + 1) % NSTAGE) {
        const long long n = std::min(chunk, N - n0);
        drain(b);
        cudaStream_t st = m_copyStream[b];
        double* h = m_pin[b].as<double>();
        double* d = m_stage[b].as<double>();
        std::memcpy(h, T + n0, sizeof(double) * n);
        std::memcpy(h + chunk, P + n0, sizeof(double) * n);
        hostCopyRows(h + 2 * chunk, K, Y + n0 * ldY, ldY, K, n);
        if (n == chunk) {
            cudaCheck(cudaMemcpyAsync(d, h, sizeof(double) * inDoubles * chunk, cudaMemcpyHostToDevice, st), "H2D state");
        } else {
            cudaCheck(cudaMemcpyAsync(d, h, sizeof(double) * n, cudaMemcpyHostToDevice, st), "H2D T");
            cudaCheck(cudaMemcpyAsync(d + chunk, h + chunk, sizeof(double) * n, cudaMemcpyHostToDevice, st), "H2D P");
            cudaCheck(cudaMemcpyAsync(d + 2 * chunk, h + 2 * chunk, sizeof(double) * n * K, cudaMemcpyHostToDevice, st), "H2D Y");
        }
        RatesArgs a = {};
        a.N = n; a.T = d; a.P = d + chunk; a.Y = d + 2 * chunk; a.ldY = K; a.ldW = K; a.ldR = R;
        double* dOut = d + (size_t) chunk * inDoubles;
        double* o = dOut;
        double** slotsW[3] = {&a.wdot, &a.cdot, &a.ddot};
        double** slotsR[5] = {&a.ropf, &a.ropr, &a.ropnet, &a.kf, &a.kc};
        for (int i = 0; i < 3; i++) if (outsW[i]) { *slotsW[i] = o; o += chunk * K; }
        for (int i = 0; i < 5; i++) if (outsR[i]) { *slotsR[i] = o; o += chunk * R; }
        launchRates(m_dev, a, full, st);
        cudaCheck(cudaGetLastError(), "ratesKernel launch");
        double* hOut = h + (size_t) chunk * inDoubles;
        if (n == chunk || nW + nR == 1) {
            const size_t cnt = n == chunk ? outDoubles * chunk : (size_t) n * (nW ? K : R);
            cudaCheck(cudaMemcpyAsync(hOut, dOut, sizeof(double) * cnt, cudaMemcpyDeviceToHost, st), "D2H");
        } else {
            size_t off = 0;
            for (int i = 0; i < 3; i++) if (outsW[i]) {
                cudaCheck(cudaMemcpyAsync(hOut + off, dOut + off, sizeof(double) * n * K, cudaMemcpyDeviceToHost, st), "D2H");
                off += (size_t) chunk * K;
            }
            for (int i = 0; i < 5; i++) if (outsR[i]) {
                cudaCheck(cudaMemcpyAsync(hOut + off, dOut + off, sizeof(double) * n * R, cudaMemcpyDeviceToHost, st), "D2H");
                off += (size_t) chunk * R;
            }
        }
        pendN0[b] = n0;
        pendN[b] = n;
    }
    for (int i = 0; i < NSTAGE; i++, b = (b + 1) % NSTAGE) drain(b);
}

void BatchedGasKinetics::quantity(int what, long long N, const double* T, const double* P, const double* Y, long long ldY,
                                  double* out, long long ld, int memspace)
{
    const int K = nSpecies(), R = nReactions();
    if (what < 0 || what >= CTB_Q_COUNT) throw CtbError("BatchedGasKinetics::quantity", "unknown quantity " + std::to_string(what));
    if (N < 0) throw CtbError("BatchedGasKinetics::quantity", "negative number of states");
    if (N == 0) return;
    if (!T || !P || !Y || !out) throw CtbError("BatchedGasKinetics::quantity", "null pointer");
    if (ldY < K) throw CtbError("BatchedGasKinetics::quantity", "ldY is smaller than the number of species");
    const bool species = what >= CTB_Q_CDOT_DDT && what <= CTB_Q_WDOT_DDC;
    const int width = species ? K : R;
    if (ld < width) throw CtbError("BatchedGasKinetics::quantity", "leading dimension " + std::to_string(ld)
                                   + " is smaller than the row width " + std::to_string(width));
    cudaCheck(cudaSetDevice(m_dev.device), "cudaSetDevice");
    DerivArgs a = {};
    a.rtolDelta = m_rtolDelta; a.skipThirdBodies = m_skip3b; a.skipFalloff = m_skipFalloff;
    auto slot = [&](DerivArgs& d) -> double** {
        if (what < CTB_Q_CDOT_DDT) { // reaction quantities: kf, ropf, ropr, ropnet x (T, P, C)
            double** fam[4] = {d.kf, d.ropf, d.ropr, d.ropnet};
            return &fam[what / 3][what % 3];
        }
        if (species) {
            double** fam[3] = {d.cdot, d.ddot, d.wdot};
            return &fam[(what - CTB_Q_CDOT_DDT) / 3][(what - CTB_Q_CDOT_DDT) % 3];
        }
        return &d.delta[what - CTB_Q_DELTA_GIBBS];
    };
    if (memspace == CTB_DEVICE) {
        a.N = N; a.T = T; a.P = P; a.Y = Y; a.ldY = ldY; a.ldR = ld; a.ldW = ld;
        *slot(a) = out;
        launchDerivs(m_dev, a, stream());
        cudaCheck(cudaGetLastError(), "derivsKernel launch");
        return;
    }
    if (memspace != CTB_HOST) throw CtbError("BatchedGasKinetics::quantity", "memspace must be CTB_HOST or CTB_DEVICE");
    const size_t perState = sizeof(double) * (2 + (size_t) K + width);
    const long long chunk = std::max<long long>(1024, std::min<long long>(N, (long long) ((64u << 20) / perState)));
    m_stage[0].reserve((size_t) chunk * perState);
    cudaStream_t st = m_copyStream[0];
    for (long long n0 = 0; n0 < N; n0 += chunk) {
        const long long n = std::min(chunk, N - n0);
        double* d = m_stage[0].as<double>();
        double* dT = d; d += chunk;
        double* dP = d; d += chunk;
        double* dY = d; d += chunk * K;
        cudaCheck(cudaMemcpyAsync(dT, T + n0, sizeof(double) * n, cudaMemcpyHostToDevice, st), "H2D T");
        cudaCheck(cudaMemcpyAsync(dP, P + n0, sizeof(double) * n, cudaMemcpyHostToDevice, st), "H2D P");
        copyRows(dY, K, Y + n0 * ldY, ldY, K, n, cudaMemcpyHostToDevice, st, "H2D Y");
        a.N = n; a.T = dT; a.P = dP; a.Y = dY; a.ldY = K; a.ldR = width; a.ldW = width;
        *slot(a) = d;
        launchDerivs(m_dev, a, st);
        cudaCheck(cudaGetLastError(), "derivsKernel launch");
        copyRows(out + n0 * ld, ld, d, width, width, n, cudaMemcpyDeviceToHost, st, "D2H");
        cudaCheck(cudaStreamSynchronize(st), "cudaStreamSynchronize");
    }
}

void BatchedGasKinetics::jacobian(long long N, const double* T, const double* P, const double* Y,
                                  long long ldY, double* jac, long long bufLen, bool ddCi, int memspace)
{
    const int K = nSpecies();
    if (N < 0) throw CtbError("BatchedGasKinetics::jacobian", "negative number of states");
    if (N == 0) return;
    if (!T || !P || !Y || !jac) throw CtbError("BatchedGasKinetics::jacobian", "null pointer");
    if (ldY < K) throw CtbError("BatchedGasKinetics::jacobian", "ldY is smaller than the number of species");
    if (bufLen < N * (long long) K * K) {
        // ArraySizeError (sourcegen/clib/templates.yaml "size_check")
        throw CtbError("BatchedGasKinetics::jacobian", "array size (" + std::to_string(bufLen)
                       + ") is too small; needs " + std::to_string(N * (long long) K * K));
    }
    cudaCheck(cudaSetDevice(m_dev.device), "cudaSetDevice");
    JacArgs a = {};
    a.ddCi = ddCi; a.skipThirdBodies = m_skip3b; a.skipFalloff = m_skipFalloff;
    if (memspace == CTB_DEVICE) {
        a.N = N; a.T = T; a.P = P; a.Y = Y; a.ldY = ldY; a.jac = jac;
        m_jacScratch[NSTAGE].reserve(jacScratchBytes(m_dev));
        launchJac(m_dev, a, stream(), m_jacScratch[NSTAGE].p);
        cudaCheck(cudaGetLastError(), "jacKernel launch");
        return;
    }
    if (memspace != CTB_HOST) throw CtbError("BatchedGasKinetics::jacobian", "memspace must be CTB_HOST or CTB_DEVICE");
    const size_t perState = sizeof(double) * (2 + (size_t) K + (size_t) K * K);
    long long chunk = std::max<long long>(256, std::min<long long>(N, (long long) ((128u << 20) / perState)));
    chunk = std::min<long long>(chunk, std::max<long long>(256, (N + 31) / 32));
    for (int b = 0; b < NSTAGE; b++) {
        m_stage[b].reserve((size_t) chunk * perState);
        m_jacScratch[b].reserve(jacScratchBytes(m_dev));
    }
    int b = 0;
    for (long long n0 = 0; n0 < N; n0 += chunk, b = (b + 1) % NSTAGE) {
        const long long n = std::min(chunk, N - n0);
        cudaStream_t st = m_copyStream[b];
        double* d = m_stage[b].as<double>();
        double* dT = d; d += chunk;
        double* dP = d; d += chunk;
        double* dY = d; d += chunk * K;
        double* dJ = d;
        cudaCheck(cudaMemcpyAsync(dT, T + n0, sizeof(double) * n, cudaMemcpyHostToDevice, st), "H2D T");
        cudaCheck(cudaMemcpyAsync(dP, P + n0, sizeof(double) * n, cudaMemcpyHostToDevice, st), "H2D P");
        copyRows(dY, K, Y + n0 * ldY, ldY, K, n, cudaMemcpyHostToDevice, st, "H2D Y");
        a.N = n; a.T = dT; a.P = dP; a.Y = dY; a.ldY = K; a.jac = dJ;
        launchJac(m_dev, a, st, m_jacScratch[b].p);
        cudaCheck(cudaGetLastError(), "jacKernel launch");
        cudaCheck(cudaMemcpyAsync(jac + n0 * (long long) K * K, dJ, sizeof(double) * n * K * K,
                                  cudaMemcpyDeviceToHost, st), "D2H jac");
    }
    for (int i = 0; i < NSTAGE; i++) cudaCheck(cudaStreamSynchronize(m_copyStream[i]), "cudaStreamSynchronize");
}

void BatchedGasKinetics::jacobianNoAlloc(long long N, const double* T, const double* P, const double* Y, long long ldY,
                                         double* jac, long long bufLen)
{
    const char* where = "BatchedGasKinetics::jacobianNoAlloc";
    const int K = nSpecies();
    if (N < 0) throw CtbError(where, "negative number of states");
    if (N > (long long) CTB_NOALLOC_MAX_STATES) {
        throw CtbError(where, std::to_string(N) + " states exceed the maximum of " + std::to_string(CTB_NOALLOC_MAX_STATES)
                       + " per call (CTB_NOALLOC_MAX_STATES); split the batch or use ctb_kin_getNetProductionRates_ddX");
    }
    if (N == 0) return;
    if (!T || !P || !Y || !jac) throw CtbError(where, "null pointer");
    if (ldY < K) throw CtbError(where, "ldY is smaller than the number of species");
    if (bufLen < N * (long long) K * K) {
        throw CtbError(where, "array size (" + std::to_string(bufLen) + ") is too small; needs "
                       + std::to_string(N * (long long) K * K));
    }
    cudaCheck(cudaSetDevice(m_dev.device), "cudaSetDevice");
    JacArgs a = {};
    a.ddCi = 0; a.skipThirdBodies = m_skip3b; a.skipFalloff = m_skipFalloff;
    a.N = N; a.T = T; a.P = P; a.Y = Y; a.ldY = ldY; a.jac = jac;
    try {
        launchJacStatic(m_dev, a, stream());
    } catch (std::exception& ex) {
        throw CtbError(where, ex.what());
    }
    cudaCheck(cudaGetLastError(), "jacTileKernel launch (static scratch)");
}

} // namespace ctb
