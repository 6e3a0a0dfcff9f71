// Host-buffer front end of k9 (kx_validate_hash_host): the call a binding of the
// reference makes with the Python strings it already holds (app.py:102 for the
// validator, app.py:305,312 for the cache key).  The batch is cut into row chunks;
// two slots, each with its own stream, run  H2D -> kernel -> D2H  so the copy of
// chunk k+1 overlaps the kernel and the read-back of chunk k.
#include <mutex>
#include "kx_common.cuh"

namespace {

struct Slot {
    cudaStream_t stream = nullptr;
    uint8_t *data = nullptr;  int64_t data_cap = 0;   // chunk bytes (+ san output of the same size)
    uint8_t *san = nullptr;
    int64_t *off = nullptr;   int64_t rows_cap = 0;
    uint8_t *verdict = nullptr;
    int64_t *hash = nullptr;
    int32_t *span = nullptr;
    int32_t *san_len = nullptr;
};

struct Staging {
    int device = -1;
    Slot slot[2];
};

// one staging area (and one lock) per device: calls for different GPUs neither wait for each other nor free and
// reallocate each other's buffers
constexpr int kMaxDevices = 64;
std::mutex g_mu[kMaxDevices];
Staging g_st[kMaxDevices];

void free_slot(Slot &s) {
    if (s.stream) cudaStreamDestroy(s.stream);
    cudaFree(s.data); cudaFree(s.san); cudaFree(s.off); cudaFree(s.verdict);
    cudaFree(s.hash); cudaFree(s.span); cudaFree(s.san_len);
    s = Slot();
}

int ensure(Staging &st, int device, int64_t rows, int64_t bytes, bool want_san) {
    st.device = device;
    for (auto &s : st.slot) {
        if (!s.stream) KX_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        if (s.data_cap < bytes + 16 || (want_san && !s.san)) {
            cudaFree(s.data); cudaFree(s.san); s.data = s.san = nullptr;
            s.data_cap = bytes + 16;
            KX_CUDA(cudaMalloc(&s.data, (size_t)s.data_cap));
            if (want_san) KX_CUDA(cudaMalloc(&s.san, (size_t)s.data_cap));
        }
        if (s.rows_cap < rows) {
            cudaFree(s.off); cudaFree(s.verdict); cudaFree(s.hash); cudaFree(s.span); cudaFree(s.san_len);
            s.rows_cap = rows;
            KX_CUDA(cudaMalloc(&s.off, sizeof(int64_t) * (size_t)(rows + 1)));
            KX_CUDA(cudaMalloc(&s.verdict, (size_t)rows));
            KX_CUDA(cudaMalloc(&s.hash, sizeof(int64_t) * (size_t)rows));
            KX_CUDA(cudaMalloc(&s.span, sizeof(int32_t) * 2 * (size_t)rows));
            KX_CUDA(cudaMalloc(&s.san_len, sizeof(int32_t) * (size_t)rows));
        }
    }
    return KX_OK;
}

}  // namespace

extern "C" void kx_release_host_staging(void) {
    int prev = 0;
    cudaGetDevice(&prev);
    for (int d = 0; d < kMaxDevices; ++d) {
        std::lock_guard<std::mutex> lk(g_mu[d]);
        if (g_st[d].device < 0) continue;
        cudaSetDevice(d);
        for (auto &s : g_st[d].slot) free_slot(s);
        g_st[d].device = -1;
    }
    cudaSetDevice(prev);
}

extern "C" int kx_validate_hash_host(const uint8_t *h_data, const int64_t *h_off, int64_t n, uint32_t flags,
                                     uint8_t *h_verdict, int64_t *h_hash, int32_t *h_span,
                                     uint8_t *h_san, int32_t *h_san_len, int device) {
    KX_REQUIRE(n >= 0, KX_EINVAL, "kx_validate_hash_host: n = %lld is negative", (long long)n);
    KX_REQUIRE((flags & ~KX_FLAG_PARSE) == 0, KX_EINVAL, "kx_validate_hash_host: unknown flags 0x%x", flags);
    if (n == 0) return KX_OK;
    KX_REQUIRE(h_off != nullptr, KX_EINVAL, "kx_validate_hash_host: h_off is NULL");
    KX_REQUIRE(h_off[0] == 0, KX_EINVAL, "kx_validate_hash_host: h_off[0] must be 0");
    KX_REQUIRE(h_data != nullptr || h_off[n] == 0, KX_EINVAL, "kx_validate_hash_host: h_data is NULL");

    const int64_t kRows = 1 << 20;
    const int64_t nchunks = (n + kRows - 1) / kRows;
    int64_t max_bytes = 0;
    for (int64_t c = 0; c < nchunks; ++c) {
        const int64_t r0 = c * kRows, r1 = (r0 + kRows < n) ? r0 + kRows : n;
        KX_REQUIRE(h_off[r1] >= h_off[r0], KX_EINVAL, "kx_validate_hash_host: offsets decrease near row %lld", (long long)r0);
        if (h_off[r1] - h_off[r0] > max_bytes) max_bytes = h_off[r1] - h_off[r0];
    }

    KX_REQUIRE(device >= 0 && device < kMaxDevices, KX_EINVAL, "kx_validate_hash_host: device ordinal %d not supported", device);
    std::lock_guard<std::mutex> lk(g_mu[device]);
    Staging &st = g_st[device];
    int prev = 0;
    KX_CUDA(cudaGetDevice(&prev));
    // Everything that can fail from here on runs inside this lambda, so that EVERY exit passes through the code
    // below it: both slot streams are drained (no copy to or from the caller's buffers is still in flight when the
    // call returns, failed or not) and the caller's current device is restored.
    auto run = [&]() -> int {
        KX_CUDA(cudaSetDevice(device));
        int rc = ensure(st, device, n < kRows ? n : kRows, max_bytes, h_san != nullptr);
        for (int64_t c = 0; c < nchunks && rc == KX_OK; ++c) {
            Slot &s = st.slot[c & 1];
            const int64_t r0 = c * kRows, r1 = (r0 + kRows < n) ? r0 + kRows : n, rows = r1 - r0;
            const int64_t b0 = h_off[r0], bytes = h_off[r1] - b0;
            if (bytes) KX_CUDA(cudaMemcpyAsync(s.data, h_data + b0, (size_t)bytes, cudaMemcpyHostToDevice, s.stream));
            KX_CUDA(cudaMemcpyAsync(s.off, h_off + r0, sizeof(int64_t) * (size_t)(rows + 1), cudaMemcpyHostToDevice, s.stream));
            // offsets stay absolute: bias the base pointers instead of rewriting them
            rc = kx_validate_hash(s.data - b0, s.off, rows, flags, h_verdict ? s.verdict : nullptr,
                                  h_hash ? s.hash : nullptr, h_span ? s.span : nullptr,
                                  h_san ? s.san - b0 : nullptr, h_san_len ? s.san_len : nullptr, s.stream);
            if (rc != KX_OK) break;
            if (h_verdict) KX_CUDA(cudaMemcpyAsync(h_verdict + r0, s.verdict, (size_t)rows, cudaMemcpyDeviceToHost, s.stream));
            if (h_hash) KX_CUDA(cudaMemcpyAsync(h_hash + r0, s.hash, sizeof(int64_t) * (size_t)rows, cudaMemcpyDeviceToHost, s.stream));
            if (h_span) KX_CUDA(cudaMemcpyAsync(h_span + 2 * r0, s.span, sizeof(int32_t) * 2 * (size_t)rows, cudaMemcpyDeviceToHost, s.stream));
            if (h_san_len) KX_CUDA(cudaMemcpyAsync(h_san_len + r0, s.san_len, sizeof(int32_t) * (size_t)rows, cudaMemcpyDeviceToHost, s.stream));
            if (h_san && bytes) KX_CUDA(cudaMemcpyAsync(h_san + b0, s.san, (size_t)bytes, cudaMemcpyDeviceToHost, s.stream));
        }
        return rc;
    };
    const int rc = run();
    cudaError_t e0 = st.slot[0].stream ? cudaStreamSynchronize(st.slot[0].stream) : cudaSuccess;
    cudaError_t e1 = st.slot[1].stream ? cudaStreamSynchronize(st.slot[1].stream) : cudaSuccess;
    cudaSetDevice(prev);
    if (rc != KX_OK) return rc;
    KX_REQUIRE(e0 == cudaSuccess && e1 == cudaSuccess, KX_ECUDA, "kx_validate_hash_host: %s",
               cudaGetErrorString(e0 != cudaSuccess ? e0 : e1));
    return KX_OK;
}
