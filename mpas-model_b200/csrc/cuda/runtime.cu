// Library lifecycle and the host-array <-> device-mirror registry behind the drop-in C ABI
// (SURVEY.md section 8b "Lifecycle"; no reference analogue -- the reference owns its arrays through
// mpas_pool_type, src/framework/mpas_pool_routines.F:4282, and never leaves the host).
// A Fortran host registers each pool array once (mpas_b200_register_field), uploads it, then calls
// the *_work entry points with the SAME host addresses it always passed; every entry point maps
// them to the device mirrors.  Device pointers pass through untouched, so the device-resident
// dycore driver (dycore.cu) uses the identical entry points.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"

#include <cstdlib>


namespace mpasb200 {

std::string g_last_error;  // message behind a non-zero return code (mpas_b200_last_error)

namespace {
struct Mirror {
    size_t bytes;
    void* dev;
};
std::map<uintptr_t, Mirror> g_reg;  // host base address -> mirror
std::mutex g_mu;
cudaStream_t g_stream = nullptr;
bool g_own_stream = false;
std::atomic<long long> g_launches{0};  // the per-block threads of the monotonic transport count too
}  // namespace

cudaStream_t stream() {
    if (!g_stream) {
        cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking);
        g_own_stream = true;
    }
    return g_stream;
}
void set_stream(cudaStream_t s) {
    if (g_own_stream && g_stream) cudaStreamDestroy(g_stream);
    g_stream = s;
    g_own_stream = false;
}
namespace {
std::mutex g_err_mu;
}
int fail(const char* where, cudaError_t e) {
    std::lock_guard<std::mutex> lk(g_err_mu);
    g_last_error = std::string(where) + ": " + cudaGetErrorString(e);
    return 1000 + (int)e;
}
int fail_msg(const std::string& msg) {
    std::lock_guard<std::mutex> lk(g_err_mu);
    g_last_error = msg;
    return 1;
}
long long launches() { return g_launches.load(); }
void count_launch(int n) { g_launches.fetch_add(n); }

// Library workspace keyed by a device address that identifies the block (an array of that block the routine is handed
// anyway).  It replaces the reference's per-phase allocate/deallocate (TI:592-626, 1029-1099) for scratch the reference
// argument lists have no room for: the three per-edge flux arrays of dyn_tend and the num_scalars-wide scratch of the
// all-scalars monotonic transport.  Keyed, not global: captured step graphs of one Dycore bake these addresses in, so a
// second (larger) Dycore in the same process must never make the first one's workspace move; several blocks of one
// process run the mono routine concurrently (one host thread per block).  A device-resident Dycore provides the storage
// itself (provide_keyed_workspace: a view of its scratch arena); for a host that calls the work routines directly the
// library allocates on first use -- never inside a stream capture.
namespace {
struct Ws {
    real* p = nullptr;
    size_t count = 0;
    bool owned = true;
};
std::mutex g_kws_mu;
std::map<std::pair<const void*, int>, Ws> g_kws;
}  // namespace
real* keyed_workspace(const void* key, size_t count, int slot) {
    std::lock_guard<std::mutex> lk(g_kws_mu);
    Ws& w = g_kws[{key, slot}];
    if (w.count < count) {
        if (!w.owned) return nullptr;  // the provider sized it; it does not grow behind the provider's back
        cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream(), &st) == cudaSuccess && st != cudaStreamCaptureStatusNone) return nullptr;  // never allocate inside a capture
        if (w.p) {
            cudaStreamSynchronize(stream());
            cudaFree(w.p);
            w.p = nullptr;
            w.count = 0;
        }
        if (cudaMalloc(&w.p, count * sizeof(real) + 64) != cudaSuccess) {
            cudaGetLastError();  // clear the sticky-free allocation error: the caller reports it
            w.p = nullptr;
            w.count = 0;
            return nullptr;
        }
        cudaMemsetAsync(w.p, 0, count * sizeof(real) + 64, stream());
        w.count = count;
    }
    return w.p;
}
void provide_keyed_workspace(const void* key, int slot, real* p, size_t count) {
    std::lock_guard<std::mutex> lk(g_kws_mu);
    Ws& w = g_kws[{key, slot}];
    if (w.p && w.owned) cudaFree(w.p);
    w.p = p;
    w.count = count;
    w.owned = false;
}
void release_keyed_workspace(const void* key) {
    std::lock_guard<std::mutex> lk(g_kws_mu);
    for (auto it = g_kws.begin(); it != g_kws.end();) {
        if (it->first.first == key) {
            if (it->second.p && it->second.owned) cudaFree(it->second.p);
            it = g_kws.erase(it);
        } else {
            ++it;
        }
    }
}

namespace {
bool g_ktime = false;
struct KRec {
    const char* name;
    cudaEvent_t e0, e1;
};
std::vector<KRec> g_krecs;
std::vector<cudaEvent_t> g_kpool;
std::map<std::string, std::pair<double, long>> g_ktot;
const char* g_kcur = nullptr;
cudaEvent_t g_ke0 = nullptr;
cudaEvent_t kev() {
    if (!g_kpool.empty()) { cudaEvent_t e = g_kpool.back(); g_kpool.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}
void kflush() {
    if (g_krecs.empty()) return;
    cudaStreamSynchronize(stream());
    for (auto& r : g_krecs) {
        float ms = 0;
        cudaEventElapsedTime(&ms, r.e0, r.e1);
        auto& t = g_ktot[r.name];
        t.first += ms;
        t.second += 1;
        g_kpool.push_back(r.e0);
        g_kpool.push_back(r.e1);
    }
    g_krecs.clear();
}
}  // namespace
bool ensure_dyn_smem(const void* kernel, size_t bytes) {
    if (bytes <= 32 * 1024) return true;  // static shared memory counts against the default 48 KB as well
    if (bytes > 220 * 1024) return false;
    static std::map<const void*, size_t> have;
    auto it = have.find(kernel);
    if (it != have.end() && it->second >= bytes) return true;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    have[kernel] = bytes;
    return true;
}

int tune(const char* name, int def) {
    std::string key = std::string("MPAS_B200_TUNE_") + name;
    const char* v = getenv(key.c_str());
    return v ? atoi(v) : def;
}

void ktime_begin(const char* name) {
    if (!g_ktime) return;
    cudaStreamCaptureStatus st;
    if (cudaStreamIsCapturing(stream(), &st) != cudaSuccess || st != cudaStreamCaptureStatusNone) return;
    g_kcur = name;
    g_ke0 = kev();
    cudaEventRecord(g_ke0, stream());
}
void ktime_end() {
    if (!g_ktime || !g_kcur) return;
    cudaEvent_t e1 = kev();
    cudaEventRecord(e1, stream());
    g_krecs.push_back({g_kcur, g_ke0, e1});
    g_kcur = nullptr;
    if (g_krecs.size() > 4096) kflush();
}

void* xlate(const void* p) {
    if (!p || g_reg.empty()) return (void*)p;
    uintptr_t a = (uintptr_t)p;
    auto it = g_reg.upper_bound(a);
    if (it == g_reg.begin()) return (void*)p;
    --it;
    if (a < it->first + it->second.bytes) return (char*)it->second.dev + (a - it->first);
    return (void*)p;
}

}  // namespace mpasb200

using namespace mpasb200;

extern "C" {

// Select the CUDA device of this rank (one process per GPU) and create the library stream.
int mpas_b200_set_device(int device) {
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail("mpas_b200_set_device", e);
    stream();
    return 0;
}

int mpas_b200_device_count() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// Register a host array (any pool array of SURVEY.md Appendix D): allocates its device mirror.
int mpas_b200_register_field(const void* host, size_t bytes) {
    std::lock_guard<std::mutex> lk(g_mu);
    uintptr_t a = (uintptr_t)host;
    auto it = g_reg.find(a);
    if (it != g_reg.end()) {
        if (it->second.bytes == bytes) return 0;
        cudaFree(it->second.dev);
        g_reg.erase(it);
    }
    Mirror m;
    m.bytes = bytes;
    cudaError_t e = cudaMalloc(&m.dev, bytes + 64);  // +64: bulk column copies read a 16-byte aligned window around a column
    if (e != cudaSuccess) return fail("mpas_b200_register_field", e);
    g_reg[a] = m;
    return 0;
}

int mpas_b200_unregister_field(const void* host) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_reg.find((uintptr_t)host);
    if (it == g_reg.end()) return fail_msg("mpas_b200_unregister_field: address not registered");
    release_edge_tile_plans(it->second.dev);
    cudaFree(it->second.dev);
    g_reg.erase(it);
    return 0;
}

// host -> device mirror (asynchronous on the library stream when the host memory is pinned)
int mpas_b200_upload(const void* host) {
    auto it = g_reg.find((uintptr_t)host);
    if (it == g_reg.end()) return fail_msg("mpas_b200_upload: address not registered");
    cudaError_t e = cudaMemcpyAsync(it->second.dev, host, it->second.bytes, cudaMemcpyHostToDevice, stream());
    if (e != cudaSuccess) return fail("mpas_b200_upload", e);
    return 0;
}

// device mirror -> host (the reference's sync point before output / restart / physics)
int mpas_b200_download(void* host) {
    auto it = g_reg.find((uintptr_t)host);
    if (it == g_reg.end()) return fail_msg("mpas_b200_download: address not registered");
    cudaError_t e = cudaMemcpyAsync(host, it->second.dev, it->second.bytes, cudaMemcpyDeviceToHost, stream());
    if (e != cudaSuccess) return fail("mpas_b200_download", e);
    return 0;
}

void* mpas_b200_device_ptr(const void* host) { return xlate(host); }

int mpas_b200_synchronize() {
    cudaError_t e = cudaStreamSynchronize(stream());
    if (e != cudaSuccess) return fail("mpas_b200_synchronize", e);
    return 0;
}

const char* mpas_b200_last_error() { return g_last_error.c_str(); }
long long mpas_b200_launch_count() { return launches(); }
int mpas_b200_real_size() { return (int)sizeof(real); }

// CUDA-event stopwatch on the library stream (bench.py times the step loop with it: torch.cuda.Event only sees
// torch's current stream)
static cudaEvent_t g_ev0 = nullptr, g_ev1 = nullptr;
int mpas_b200_timer_start() {
    if (!g_ev0) { cudaEventCreate(&g_ev0); cudaEventCreate(&g_ev1); }
    cudaError_t e = cudaEventRecord(g_ev0, stream());
    if (e != cudaSuccess) return fail("mpas_b200_timer_start", e);
    return 0;
}
double mpas_b200_timer_stop() {
    if (!g_ev0) return -1.0;
    cudaEventRecord(g_ev1, stream());
    cudaError_t e = cudaEventSynchronize(g_ev1);
    if (e != cudaSuccess) { fail("mpas_b200_timer_stop", e); return -1.0; }
    float ms = 0;
    cudaEventElapsedTime(&ms, g_ev0, g_ev1);
    return (double)ms;
}

// per-kernel timing: enable, run eager steps, then read "name ms calls" lines
int mpas_b200_ktime_enable(int on) {
    kflush();
    g_ktime = on != 0;
    if (on) g_ktot.clear();
    return 0;
}
int mpas_b200_ktime_report(char* buf, int cap) {
    kflush();
    std::string out;
    for (auto& kv : g_ktot) {
        char line[256];
        std::string nm;  // "(kernel<5, false>)" -> "kernel<5,false>": one token per line for the parser
        for (char c : kv.first)
            if (c != ' ' && c != '(' && c != ')') nm += c;
        snprintf(line, sizeof(line), "%s %.6f %ld\n", nm.c_str(), kv.second.first, kv.second.second);
        out += line;
    }
    if ((int)out.size() + 1 > cap) return (int)out.size() + 1;
    std::memcpy(buf, out.c_str(), out.size() + 1);
    return 0;
}

int mpas_b200_finalize() {
    std::lock_guard<std::mutex> lk(g_mu);
    release_edge_tile_plans(nullptr);
    for (auto& kv : g_reg) cudaFree(kv.second.dev);
    g_reg.clear();
    {
        std::lock_guard<std::mutex> lk2(g_kws_mu);
        for (auto& kv : g_kws)
            if (kv.second.p && kv.second.owned) cudaFree(kv.second.p);
        g_kws.clear();
    }
    return 0;
}

}  // extern "C"
