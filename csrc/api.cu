// Host-side runtime pieces of libdinunet_b200: error reporting, device-memory arena +
// CUDA-IPC peer mapping (the symmetric-memory substrate the fused aggregation kernels use),
// TMA descriptor encoding, and the threaded CPU reducer used by the COINSTAC file mode.
#include "common.cuh"

#include <stdarg.h>
#include <stdlib.h>

#include <atomic>
#include <mutex>
#include <thread>
#include <vector>

#include "../include/dinunet_b200.h"

// ------------------------------------------------------------------------------------ errors
static thread_local char g_err[1024] = {0};
static std::atomic<long long> g_launches{0};

void dnb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

DNB_EXPORT const char* dnb_last_error(void) { return g_err; }
DNB_EXPORT int dnb_version(void) { return 100; }

void dnb_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

bool dnb_pdl_enabled() {
    static const bool on = [] {
        const char* e = getenv("DNB_PDL");
        return !(e && e[0] == '0');   // on unless DNB_PDL=0: see common.cuh
    }();
    return on;
}
DNB_EXPORT long long dnb_launch_count(void) { return g_launches.load(); }
DNB_EXPORT void dnb_reset_launch_count(void) { g_launches.store(0); }

// ------------------------------------------------------------------------------------ arena / IPC
// The arena is plain cudaMalloc memory so that cudaIpcGetMemHandle describes exactly this
// allocation (torch's caching allocator sub-allocates, which would need base+offset games).
DNB_EXPORT int dnb_arena_alloc(size_t bytes, void** out) {
    DNB_REQUIRE(out != nullptr, "dnb_arena_alloc: out is null");
    DNB_CUDA_CHECK(cudaMalloc(out, bytes));
    DNB_CUDA_CHECK(cudaMemset(*out, 0, bytes));
    DNB_CUDA_CHECK(cudaDeviceSynchronize());
    return 0;
}
DNB_EXPORT int dnb_arena_free(void* p) {
    DNB_CUDA_CHECK(cudaFree(p));
    return 0;
}
DNB_EXPORT int dnb_ipc_handle_size(void) { return (int)sizeof(cudaIpcMemHandle_t); }
DNB_EXPORT int dnb_ipc_get_handle(void* dptr, void* handle_out) {
    cudaIpcMemHandle_t h;
    DNB_CUDA_CHECK(cudaIpcGetMemHandle(&h, dptr));
    memcpy(handle_out, &h, sizeof(h));
    return 0;
}
DNB_EXPORT int dnb_ipc_open_handle(const void* handle, void** out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    DNB_CUDA_CHECK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
DNB_EXPORT int dnb_ipc_close_handle(void* p) {
    DNB_CUDA_CHECK(cudaIpcCloseMemHandle(p));
    return 0;
}
DNB_EXPORT int dnb_enable_peer_access(int peer_device) {
    int can = 0, dev = 0;
    DNB_CUDA_CHECK(cudaGetDevice(&dev));
    if (peer_device == dev) return 0;
    DNB_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, dev, peer_device));
    DNB_REQUIRE(can, "device %d cannot access peer %d", dev, peer_device);
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        return 0;
    }
    DNB_CUDA_CHECK(e);
    return 0;
}

DNB_EXPORT int dnb_copy_async(void* dst, const void* src, size_t bytes, int kind, void* stream_) {
    DNB_REQUIRE(dst && src && kind >= 1 && kind <= 3, "dnb_copy_async: bad arguments");
    const cudaMemcpyKind k = kind == 1 ? cudaMemcpyHostToDevice : (kind == 2 ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice);
    DNB_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, k, reinterpret_cast<cudaStream_t>(stream_)));
    return 0;
}

// ------------------------------------------------------------------------------------ TMA descriptors
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
    static PFN_encodeTiled fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiled)p;
    });
    return fn;
}

// Row-major bf16 matrix [rows, cols] with leading dimension ld (elements); box = [box_rows, 64 cols]
// (64 bf16 = 128 B = one swizzle-128B row).  Out-of-bounds elements read as zero.
int dnb_make_tmap_bf16_2d(CUtensorMap* map, const void* base, long rows, long cols, long ld, int box_rows,
                          int box_cols) {
    PFN_encodeTiled enc = get_encode();
    DNB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled unavailable (driver too old?)");
    DNB_REQUIRE(((uintptr_t)base & 15) == 0, "TMA base pointer must be 16-byte aligned");
    DNB_REQUIRE((ld * 2) % 16 == 0, "TMA row pitch must be a multiple of 16 bytes (ld=%ld)", ld);
    DNB_REQUIRE(box_cols * 2 <= 128 && box_rows <= 256, "TMA box too large");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DNB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed: %d (rows=%ld cols=%ld ld=%ld)", (int)r, rows, cols,
                ld);
    return 0;
}

// ------------------------------------------------------------------------------------ CPU reducer
// Mean over `n_src` same-length arrays (fp32 or fp16 wire format) into fp32 — the aggregator's
// job in COINSTAC file mode.  Replaces the reference's multiprocessing.Pool of numpy workers
// (remote.py:20-21) with threads over contiguous chunks: no pickling, one pass over memory.
static inline float half_bits_to_float(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000) << 16;
    uint32_t exp = (h >> 10) & 0x1F;
    uint32_t man = h & 0x3FF;
    uint32_t f;
    if (exp == 0) {
        if (man == 0) {
            f = sign;
        } else {
            exp = 127 - 15 + 1;
            while (!(man & 0x400)) {
                man <<= 1;
                exp--;
            }
            man &= 0x3FF;
            f = sign | (exp << 23) | (man << 13);
        }
    } else if (exp == 31) {
        f = sign | 0x7F800000u | (man << 13);
    } else {
        f = sign | ((exp + 127 - 15) << 23) | (man << 13);
    }
    float out;
    memcpy(&out, &f, 4);
    return out;
}

DNB_EXPORT int dnb_cpu_mean_f32(const void* const* srcs, int n_src, long long n, int is_half, float* out,
                                int n_threads) {
    DNB_REQUIRE(n_src > 0 && n >= 0 && out != nullptr, "dnb_cpu_mean_f32: bad arguments");
    if (n_threads <= 0) {
        n_threads = (int)std::thread::hardware_concurrency();
        if (n_threads <= 0) n_threads = 4;
        if (n_threads > 16) n_threads = 16;
    }
    if (n < (1 << 16)) n_threads = 1;
    const float inv = 1.0f / (float)n_src;
    auto work = [&](long long lo, long long hi) {
        for (long long i = lo; i < hi; ++i) {
            float acc = 0.f;
            for (int s = 0; s < n_src; ++s)
                acc += is_half ? half_bits_to_float(((const uint16_t*)srcs[s])[i]) : ((const float*)srcs[s])[i];
            out[i] = acc * inv;
        }
    };
    if (n_threads == 1) {
        work(0, n);
        return 0;
    }
    std::vector<std::thread> th;
    long long per = (n + n_threads - 1) / n_threads;
    for (int t = 0; t < n_threads; ++t) {
        long long lo = t * per, hi = lo + per > n ? n : lo + per;
        if (lo >= hi) break;
        th.emplace_back(work, lo, hi);
    }
    for (auto& t : th) t.join();
    return 0;
}

// ------------------------------------------------------------------------------------ native data loading
// One FreeSurfer aseg-stats file: a header line, then `name<TAB>value` rows (66 of them in the
// reference's test_fsl data).  Returns the number of values parsed (negative on error); with
// `normalise` every value is divided by the subject's own maximum, as the reference dataset does
// (comps/fs/__init__.py:37).  Replaces a pandas.read_csv per sample per epoch.
DNB_EXPORT int dnb_cpu_orthogonalize(float* P, long long n, int r, float eps, float rel) {
    DNB_REQUIRE(P != nullptr && n >= 0 && r >= 1 && r <= 1024, "dnb_cpu_orthogonalize: bad arguments");
    std::vector<double> orig(r, 0.0), dot(r, 0.0);
    double fro = 0.0;
    for (long long i = 0; i < n; ++i)
        for (int c = 0; c < r; ++c) {
            const double v = P[i * r + c];
            orig[c] += v * v;
        }
    for (int c = 0; c < r; ++c) { fro += orig[c]; orig[c] = std::sqrt(orig[c]); }
    const double scale = std::sqrt(fro) + eps;
    for (int c = 0; c < r; ++c) {
        double nn = 0.0;
        for (long long i = 0; i < n; ++i) nn += (double)P[i * r + c] * P[i * r + c];
        const double nrm = std::sqrt(nn);
        const double thr = std::max((double)eps * scale, (double)rel * orig[c]);
        if (nrm <= thr) {
            for (long long i = 0; i < n; ++i) P[i * r + c] = 0.f;
            continue;
        }
        const double inv = 1.0 / nrm;
        for (int j = c + 1; j < r; ++j) dot[j] = 0.0;
        for (long long i = 0; i < n; ++i) {
            float* row = P + i * r;
            const float q = (float)(row[c] * inv);
            row[c] = q;
            for (int j = c + 1; j < r; ++j) dot[j] += (double)q * row[j];
        }
        for (long long i = 0; i < n; ++i) {
            float* row = P + i * r;
            const double q = row[c];
            for (int j = c + 1; j < r; ++j) row[j] = (float)(row[j] - q * dot[j]);
        }
    }
    return 0;
}

DNB_EXPORT int dnb_parse_aseg(const char* path, double* out, int max_n, int normalise) {
    if (!path || !out || max_n <= 0) return -1;
    FILE* f = fopen(path, "r");
    if (!f) return -2;
    char line[4096];
    int n = 0;
    bool first = true;
    double mx = 0.0;
    while (fgets(line, sizeof(line), f)) {
        if (first) { first = false; continue; }
        char* tab = strchr(line, '\t');
        if (!tab) continue;
        char* end = nullptr;
        const double v = strtod(tab + 1, &end);
        if (end == tab + 1) continue;
        if (n < max_n) out[n] = v;
        if (n == 0 || v > mx) mx = v;
        ++n;
    }
    fclose(f);
    if (n > max_n) return -3;
    if (normalise && n > 0 && mx != 0.0)
        for (int i = 0; i < n; ++i) out[i] /= mx;
    return n;
}
