// Library runtime: error state, launch accounting, device queries, tensor-map cache.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace igg {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

// ---- tensor maps -------------------------------------------------------------------------------
// cuTensorMapEncodeTiled is fetched through the runtime so the library does not link libcuda.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    });
    return fn;
}

// L2 fetch granularity of TMA loads (IGG_L2_PROMO = 0 none, 1 64 B, 2 128 B, 3 256 B)
static CUtensorMapL2promotion l2_promotion() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("IGG_L2_PROMO");
        v = e ? atoi(e) : 1;
        if (v < 0 || v > 3) v = 1;
    }
    return (CUtensorMapL2promotion)v;
}

struct MapKey {
    const void* base;
    int rank;
    uint64_t dims[3];
    uint32_t box[3];
    bool operator==(const MapKey& o) const { return memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        const uint64_t* w = (const uint64_t*)&k;
        size_t h = 1469598103934665603ull;
        for (size_t i = 0; i < sizeof(MapKey) / 8; ++i) h = (h ^ w[i]) * 1099511628211ull;
        return h;
    }
};

int get_tensor_map_f64(const void* base, int rank, const uint64_t* dims, const uint32_t* box, CUtensorMap* out) {
    static std::mutex mu;
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    IGG_REQUIRE(rank >= 2 && rank <= 3, "tensor map rank must be 2 or 3");
    MapKey key;
    memset(&key, 0, sizeof(key));
    key.base = base;
    key.rank = rank;
    for (int i = 0; i < rank; ++i) {
        key.dims[i] = dims[i];
        key.box[i] = box[i];
    }
    // A time loop asks for the same two to six maps on every launch: a small per-thread table of the most recent
    // ones answers without hashing or taking the mutex (launch preparation is what bounds small-grid steps).
    constexpr int RECENT = 8;
    static thread_local MapKey recent_key[RECENT];
    static thread_local CUtensorMap recent_map[RECENT];
    static thread_local int recent_n = 0, recent_next = 0;
    for (int i = 0; i < recent_n; ++i)
        if (memcmp(&recent_key[i], &key, sizeof(MapKey)) == 0) {
            *out = recent_map[i];
            return 0;
        }
    auto remember = [&](const CUtensorMap& m) {
        recent_key[recent_next] = key;
        recent_map[recent_next] = m;
        recent_next = (recent_next + 1) % RECENT;
        if (recent_n < RECENT) ++recent_n;
    };
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(key);
        if (it != cache.end()) {
            *out = it->second;
            remember(it->second);
            return 0;
        }
    }
    EncodeTiledFn fn = encode_fn();
    IGG_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
    IGG_REQUIRE(((uintptr_t)base & 15) == 0, "TMA needs a 16-byte aligned base address");
    cuuint64_t gdims[3];
    cuuint64_t gstrides[2];
    cuuint32_t gbox[3], estr[3];
    for (int i = 0; i < rank; ++i) {
        gdims[i] = dims[i];
        gbox[i] = box[i];
        estr[i] = 1;
        IGG_REQUIRE(box[i] >= 1 && box[i] <= 256, "TMA box dimension out of range (1..256)");
    }
    uint64_t s = dims[0] * sizeof(double);
    for (int i = 0; i < rank - 1; ++i) {
        gstrides[i] = s;
        IGG_REQUIRE((s & 15) == 0, "TMA needs row pitches that are multiples of 16 bytes (even inner extent)");
        s *= dims[i + 1];
    }
    IGG_REQUIRE((box[0] * sizeof(double)) % 16 == 0, "TMA inner box must be a multiple of 16 bytes");
    CUtensorMap m;
    CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, (cuuint32_t)rank, const_cast<void*>(base), gdims,
                    gstrides, gbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                    l2_promotion(), CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
        return IGG_ERR_UNSUPPORTED;
    }
    {
        std::lock_guard<std::mutex> lk(mu);
        if (cache.size() > 4096) cache.clear();
        cache[key] = m;
    }
    remember(m);
    *out = m;
    return 0;
}

}  // namespace igg

extern "C" {

int igg_version(void) { return 100; }

const char* igg_last_error(void) { return igg::g_err; }

unsigned long long igg_launch_count(void) { return igg::g_launches.load(std::memory_order_relaxed); }

int igg_device_info(int* sms, int* cc_major, int* cc_minor, size_t* smem_optin) {
    int dev = 0;
    IGG_CUDA_OK(cudaGetDevice(&dev));
    int v = 0;
    if (sms) {
        IGG_CUDA_OK(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev));
        *sms = v;
    }
    if (cc_major) {
        IGG_CUDA_OK(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, dev));
        *cc_major = v;
    }
    if (cc_minor) {
        IGG_CUDA_OK(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, dev));
        *cc_minor = v;
    }
    if (smem_optin) {
        IGG_CUDA_OK(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        *smem_optin = (size_t)v;
    }
    return 0;
}

}  // extern "C"
