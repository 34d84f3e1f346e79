// Device-resident dycore driver: the atm_srk3 call sequence (TI:142-1796) over fields that live in HBM.
#pragma once
#include <atomic>
#include <condition_variable>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "halo.h"

namespace mpasb200 {

struct DField {
    void* dev = nullptr;
    size_t bytes = 0;
    int kind = 0;   // 0 real, 1 int
    int loc = 0;    // 0 none, 1 cell, 2 edge, 3 vertex
    int inner = 1;  // product of all dims but the last
    bool owned = true;  // false: a view of library workspace (not freed with the block)
    std::vector<long long> dims;
};

// A view of the block's scratch arena: phase-local work arrays of different phases share storage (the reference
// allocates and frees them around each phase for the same reason, TI:592-626, 1029-1099).
struct ArenaView {
    std::string name;
    int group;  // 1 = dyn_tend rk 1 scratch, 2 = monotonic transport scratch, 3 = advance_scalars flux
    real* garbage;      // first element of the trailing "garbage" column (zero-filled at the start of the phase, as TI:593-605)
    int inner;
};

struct DBlock {
    std::map<std::string, DField> f;
    char* arena = nullptr;
    size_t arena_bytes = 0;
    std::vector<ArenaView> views;
    std::map<std::string, double> cfg;
    HaloBlock halo;
    real* R(const std::string& n);
    real* Ropt(const std::string& n);
    int* I(const std::string& n);
    double c(const std::string& n) const;
    double c(const std::string& n, double dflt) const;
    void swap_levels(const std::string& n) { std::swap(f[n + "@1"].dev, f[n + "@2"].dev); }
};

struct PhaseTimer {
    double ms = 0;
    long calls = 0;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;
};

struct HostBarrier {  // barrier of the per-block host threads of the monotonic transport; abortable: a thread that fails
                      // before it reaches the barrier must not leave the others blocked forever
    std::mutex m;
    std::condition_variable cv;
    int count = 0, gen = 0, n = 1;
    bool aborted = false;
    void reset(int nthreads) {
        std::lock_guard<std::mutex> lk(m);
        n = nthreads; count = 0; aborted = false;
    }
    void abort() {
        std::lock_guard<std::mutex> lk(m);
        aborted = true;
        cv.notify_all();
    }
    bool wait() {  // false: aborted
        std::unique_lock<std::mutex> lk(m);
        if (aborted) return false;
        int g = gen;
        if (++count == n) { count = 0; ++gen; cv.notify_all(); }
        else cv.wait(lk, [&] { return g != gen || aborted; });
        return !aborted;
    }
};

struct Dycore {
    std::vector<DBlock> blocks;
    int rank = 0, nranks = 1;
    Nccl nccl;
    bool have_comm = false;
    P2P p2p;  // peer-to-peer windows, set up after the second (eager) step when every rank can map every other
    long nExchanges = 0;
    bool timing = false;
    bool aggregate = true;
    bool fold_first = true;  // first acoustic substep's edge pass folded into the divergence damping (srk3.cu, acoustic.cu); option "fold_first"
    bool use_graph = false;
    bool diet = true;  // phase-local scratch shares one arena; optional (physics / output-only) arrays get no mirror
    std::map<std::string, PhaseTimer> timers;
    std::vector<std::string> timer_order;
    std::vector<cudaEvent_t> event_pool;
    HostBarrier bar;
    std::string err;
    std::mutex err_mu;
    // captured step graph
    cudaGraphExec_t graph_exec[2] = {nullptr, nullptr};
    long long graph_nodes[2] = {0, 0};
    long graph_exchanges[2] = {0, 0};
    long steps_done = 0;
    int parity = 0;  // number of steps taken mod 2 (time levels swap every step)
    real* d_reduce = nullptr;
    real* d_minmax = nullptr;
};

struct XField {
    std::string name;
    int mask;  // halo layer mask, 0 = all
};

int dycore_exchange(Dycore& m, const std::vector<XField>& fields);
int dycore_p2p_setup(Dycore& m);   // collective; leaves m.p2p.on false (NCCL path) when any rank cannot
void dycore_p2p_teardown(Dycore& m);
int dycore_check_p2p(Dycore& m);   // non-zero (with a message) once a peer-to-peer wait has given up
int dycore_init(Dycore& m);
int dycore_step(Dycore& m);
int dycore_call(Dycore& m, const std::string& routine, const double* prm, int nprm);
int dycore_zero_garbage(DBlock& b, int group);  // zero the garbage columns of the arena views of one phase

}  // namespace mpasb200
