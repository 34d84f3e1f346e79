// Device-resident dycore: field store, halo-exchange groups, C ABI (mpas_b200_dycore_*).
// The step sequence itself is in srk3.cu.
#include "dycore.h"

#include <cstring>
#include <stdexcept>

#include "../host/decomp.h"

namespace mpas {
extern std::string g_last_error;
}

namespace mpasb200 {

real* DBlock::R(const std::string& n) {
    auto it = f.find(n);
    if (it == f.end()) throw std::runtime_error("dycore: missing field " + n);
    return (real*)it->second.dev;
}
real* DBlock::Ropt(const std::string& n) {
    auto it = f.find(n);
    return it == f.end() ? nullptr : (real*)it->second.dev;
}
int* DBlock::I(const std::string& n) {
    auto it = f.find(n);
    if (it == f.end()) throw std::runtime_error("dycore: missing field " + n);
    return (int*)it->second.dev;
}
double DBlock::c(const std::string& n) const {
    auto it = cfg.find(n);
    if (it == cfg.end()) throw std::runtime_error("dycore: missing config " + n);
    return it->second;
}
double DBlock::c(const std::string& n, double dflt) const {
    auto it = cfg.find(n);
    return it == cfg.end() ? dflt : it->second;
}

// One exchange of a group of fields (adjacent reference exchanges with no compute in between are
// aggregated into one message per neighbour).
int dycore_exchange(Dycore& m, const std::vector<XField>& fields) {
    m.nExchanges += (long)fields.size();
    const int P = (int)m.blocks.size();
    if (P == 1 && m.nranks == 1) return 0;
    std::string key;
    for (const XField& xf : fields) key += xf.name + ":" + std::to_string(xf.mask) + ";";
    std::vector<HaloPlan*> plans(P);
    for (int b = 0; b < P; ++b) {
        DBlock& blk = m.blocks[b];
        // the plan caches array addresses; time levels swap every step, so the key carries the address
        std::string k2 = key;
        for (const XField& xf : fields) k2 += std::to_string((uintptr_t)blk.R(xf.name)) + ",";
        auto it = blk.halo.plans.find(k2);
        if (it == blk.halo.plans.end()) {
            std::vector<HaloItem> items;
            for (const XField& xf : fields) {
                const DField& df = blk.f.at(xf.name);
                if (df.loc < 1 || df.loc > 3) throw std::runtime_error("dycore_exchange: field " + xf.name + " is not decomposed");
                HaloItem hi;
                hi.array = (real*)df.dev; hi.loc = df.loc; hi.inner = df.inner; hi.mask = xf.mask;
                items.push_back(hi);
            }
            HaloPlan* p = halo_build_plan(blk.halo, items);
            blk.halo.plans[k2] = p;
            plans[b] = p;
        } else {
            plans[b] = it->second;
        }
    }
    for (int b = 0; b < P; ++b) {
        int rc = halo_pack(plans[b]);
        if (rc) return rc;
    }
    if (m.nranks > 1) {
        if (!m.have_comm) return fail_msg("dycore_exchange: nranks > 1 but no NCCL communicator (call mpas_b200_dycore_comm_init)");
        HaloPlan* p = plans[0];
        if (m.nccl.group_start()) return fail_msg(std::string("ncclGroupStart: ") + m.nccl.last_error());
        for (size_t i = 0; i < p->recv.nbr.size(); ++i)
            if (m.nccl.recv(p->recv.d_buf + p->recv.off[i], (size_t)p->recv.cnt[i], p->recv.nbr[i], stream())) return fail_msg(std::string("ncclRecv: ") + m.nccl.last_error());
        for (size_t i = 0; i < p->send.nbr.size(); ++i)
            if (m.nccl.send(p->send.d_buf + p->send.off[i], (size_t)p->send.cnt[i], p->send.nbr[i], stream())) return fail_msg(std::string("ncclSend: ") + m.nccl.last_error());
        if (m.nccl.group_end()) return fail_msg(std::string("ncclGroupEnd: ") + m.nccl.last_error());
        count_launch();
    } else {
        // blocks of one process: message = device copy from the sender's packed segment to the receiver's
        for (int b = 0; b < P; ++b) {
            HaloPlan* ps = plans[b];
            for (size_t i = 0; i < ps->send.nbr.size(); ++i) {
                const int dstBlock = ps->send.nbr[i];
                if (dstBlock < 0 || dstBlock >= P) return fail_msg("dycore_exchange: neighbour block not in this process");
                HaloPlan* pr = plans[dstBlock];
                size_t j = 0;
                for (; j < pr->recv.nbr.size(); ++j)
                    if (pr->recv.nbr[j] == m.blocks[b].halo.blockID) break;
                if (j == pr->recv.nbr.size() || pr->recv.cnt[j] != ps->send.cnt[i]) return fail_msg("dycore_exchange: unmatched exchange lists");
                cudaError_t e = cudaMemcpyAsync(pr->recv.d_buf + pr->recv.off[j], ps->send.d_buf + ps->send.off[i], (size_t)ps->send.cnt[i] * sizeof(real),
                                                cudaMemcpyDeviceToDevice, stream());
                if (e != cudaSuccess) return fail("dycore_exchange copy", e);
                count_launch();
            }
        }
    }
    for (int b = 0; b < P; ++b) {
        int rc = halo_unpack(plans[b]);
        if (rc) return rc;
    }
    return 0;
}

namespace {

int add_field(DBlock& blk, const std::string& name, const void* host, int kind, int loc, const std::vector<long long>& dims) {
    DField df;
    df.kind = kind; df.loc = loc; df.dims = dims;
    size_t n = 1;
    for (auto d : dims) n *= (size_t)d;
    df.inner = 1;
    for (size_t k = 0; k + 1 < dims.size(); ++k) df.inner *= (int)dims[k];
    df.bytes = n * (kind == 0 ? sizeof(real) : sizeof(int));
    cudaError_t e = cudaMalloc(&df.dev, df.bytes + 64);  // +64: bulk column copies read a 16-byte aligned window around a column
    if (e != cudaSuccess) return fail(("cudaMalloc " + name).c_str(), e);
    if (host) {
        if (kind == 0 && sizeof(real) != sizeof(double)) {
            std::vector<real> tmp(n);
            const double* src = (const double*)host;
            for (size_t i = 0; i < n; ++i) tmp[i] = (real)src[i];
            e = cudaMemcpy(df.dev, tmp.data(), df.bytes, cudaMemcpyHostToDevice);
        } else {
            e = cudaMemcpy(df.dev, host, df.bytes, cudaMemcpyHostToDevice);
        }
        if (e != cudaSuccess) return fail(("upload " + name).c_str(), e);
    } else {
        cudaMemset(df.dev, 0, df.bytes);
    }
    auto it = blk.f.find(name);
    if (it != blk.f.end() && it->second.owned) cudaFree(it->second.dev);
    blk.f[name] = df;
    return 0;
}

void copy_lists(const std::vector<mpas::ExchangeList>& in, std::vector<HaloList>& out) {
    for (const auto& l : in) {
        HaloList h;
        h.endPoint = l.endPointID; h.layer = l.haloLayer; h.src = l.srcList; h.dest = l.destList;
        out.push_back(h);
    }
}

}  // namespace
}  // namespace mpasb200

using namespace mpasb200;

#define DY_TRY try {
#define DY_CATCH                                 \
    }                                            \
    catch (const std::exception& e) {            \
        mpas::g_last_error = e.what();           \
        return 1;                                \
    }

extern "C" {

void* mpas_b200_dycore_create(int nLocalBlocks, int rank, int nranks) {
    Dycore* m = new Dycore();
    m->blocks.resize(nLocalBlocks);
    m->rank = rank;
    m->nranks = nranks;
    stream();
    return m;
}

void mpas_b200_dycore_destroy(void* h) {
    Dycore* m = (Dycore*)h;
    if (!m) return;
    cudaStreamSynchronize(stream());
    for (auto& b : m->blocks) {
        auto sa = b.f.find("scale_arr");
        if (sa != b.f.end()) release_keyed_workspace(sa->second.dev);  // the all-scalars mono workspace of this block
        for (auto& kv : b.f)
            if (kv.second.owned) cudaFree(kv.second.dev);
        for (auto& kv : b.halo.plans) halo_free_plan(kv.second);
    }
    for (int i = 0; i < 2; ++i)
        if (m->graph_exec[i]) cudaGraphExecDestroy(m->graph_exec[i]);
    for (auto e : m->event_pool) cudaEventDestroy(e);
    if (m->d_reduce) cudaFree(m->d_reduce);
    if (m->d_minmax) cudaFree(m->d_minmax);
    m->nccl.destroy();
    delete m;
}

int mpas_b200_dycore_attach_block(void* h, int iLocal, void* hostBlock) {
    DY_TRY
    Dycore* m = (Dycore*)h;
    mpas::Block* hb = (mpas::Block*)hostBlock;
    if (!hb) return fail_msg("mpas_b200_dycore_attach_block: null host block");
    DBlock& blk = m->blocks.at(iLocal);
    for (auto& kv : hb->pool.f) {
        const mpas::Field& fl = kv.second;
        std::vector<long long> dims(fl.dims.begin(), fl.dims.end());
        const void* src = fl.kind == mpas::FK_REAL ? (const void*)fl.r.data() : (const void*)fl.i.data();
        int rc = add_field(blk, kv.first, src, fl.kind == mpas::FK_REAL ? 0 : 1, (int)fl.loc, dims);
        if (rc) return rc;
    }
    for (auto& kv : hb->pool.cfg) blk.cfg[kv.first] = kv.second;
    blk.halo.blockID = hb->blockID;
    copy_lists(hb->cellExch.send, blk.halo.send[1]); copy_lists(hb->cellExch.recv, blk.halo.recv[1]);
    copy_lists(hb->edgeExch.send, blk.halo.send[2]); copy_lists(hb->edgeExch.recv, blk.halo.recv[2]);
    copy_lists(hb->vertexExch.send, blk.halo.send[3]); copy_lists(hb->vertexExch.recv, blk.halo.recv[3]);
    return 0;
    DY_CATCH
}

int mpas_b200_dycore_set_cfg(void* h, int iLocal, const char* name, double value) {
    ((Dycore*)h)->blocks.at(iLocal).cfg[name] = value;
    return 0;
}
double mpas_b200_dycore_get_cfg(void* h, int iLocal, const char* name) { return ((Dycore*)h)->blocks.at(iLocal).c(name, 0.0); }

int mpas_b200_nccl_unique_id(void* id128) {
    Nccl n;
    std::string err;
    if (!n.load(err)) return fail_msg(err);
    if (n.get_unique_id(id128)) return fail_msg(std::string("ncclGetUniqueId: ") + n.last_error());
    return 0;
}

int mpas_b200_dycore_comm_init(void* h, const void* id128) {
    Dycore* m = (Dycore*)h;
    std::string err;
    if (!m->nccl.load(err)) return fail_msg(err);
    if (m->nccl.init_rank(id128, m->nranks, m->rank)) return fail_msg(std::string("ncclCommInitRank: ") + m->nccl.last_error());
    m->have_comm = true;
    return 0;
}

int mpas_b200_dycore_init(void* h) {
    DY_TRY
    return dycore_init(*(Dycore*)h);
    DY_CATCH
}

int mpas_b200_dycore_step(void* h, int nsteps) {
    DY_TRY
    Dycore& m = *(Dycore*)h;
    for (int i = 0; i < nsteps; ++i) {
        int rc = dycore_step(m);
        if (rc) return rc;
    }
    return 0;
    DY_CATCH
}

int mpas_b200_dycore_get_field(void* h, int iLocal, const char* name, void* host_out, size_t bytes) {
    DY_TRY
    DBlock& b = ((Dycore*)h)->blocks.at(iLocal);
    auto it = b.f.find(name);
    if (it == b.f.end()) return fail_msg(std::string("no field ") + name);
    if (bytes != it->second.bytes) return fail_msg(std::string("size mismatch for field ") + name);
    cudaError_t e = cudaMemcpyAsync(host_out, it->second.dev, bytes, cudaMemcpyDeviceToHost, stream());
    if (e != cudaSuccess) return fail("mpas_b200_dycore_get_field", e);
    e = cudaStreamSynchronize(stream());
    if (e != cudaSuccess) return fail("mpas_b200_dycore_get_field sync", e);
    return 0;
    DY_CATCH
}

int mpas_b200_dycore_set_field(void* h, int iLocal, const char* name, const void* host_in, size_t bytes) {
    DY_TRY
    DBlock& b = ((Dycore*)h)->blocks.at(iLocal);
    auto it = b.f.find(name);
    if (it == b.f.end()) return fail_msg(std::string("no field ") + name);
    if (bytes != it->second.bytes) return fail_msg(std::string("size mismatch for field ") + name);
    cudaError_t e = cudaMemcpyAsync(it->second.dev, host_in, bytes, cudaMemcpyHostToDevice, stream());
    if (e != cudaSuccess) return fail("mpas_b200_dycore_set_field", e);
    return 0;
    DY_CATCH
}

void* mpas_b200_dycore_field_device_ptr(void* h, int iLocal, const char* name) {
    DBlock& b = ((Dycore*)h)->blocks.at(iLocal);
    auto it = b.f.find(name);
    return it == b.f.end() ? nullptr : it->second.dev;
}

int mpas_b200_dycore_exchange(void* h, const char* name, int layer) {
    DY_TRY
    return dycore_exchange(*(Dycore*)h, {{name, layer == 0 ? 0 : (1 << (layer - 1))}});
    DY_CATCH
}

int mpas_b200_dycore_call(void* h, const char* routine, const double* params, int nparams) {
    DY_TRY
    return dycore_call(*(Dycore*)h, routine, params, nparams);
    DY_CATCH
}

long mpas_b200_dycore_exchange_count(void* h) { return ((Dycore*)h)->nExchanges; }

int mpas_b200_dycore_set_option(void* h, const char* name, int value) {
    Dycore* m = (Dycore*)h;
    std::string n = name;
    if (n == "timing") m->timing = value != 0;
    else if (n == "graph") m->use_graph = value != 0;
    else if (n == "aggregate") m->aggregate = value != 0;
    else return fail_msg("unknown option " + n);
    return 0;
}

int mpas_b200_dycore_timer_count(void* h) { return (int)((Dycore*)h)->timer_order.size(); }
int mpas_b200_dycore_timer_get(void* h, int idx, char* name, int cap, double* ms, long* calls) {
    Dycore* m = (Dycore*)h;
    if (idx < 0 || idx >= (int)m->timer_order.size()) return 1;
    const std::string& n = m->timer_order[idx];
    std::strncpy(name, n.c_str(), cap - 1);
    name[cap - 1] = 0;
    PhaseTimer& t = m->timers[n];
    *ms = t.ms;
    *calls = t.calls;
    return 0;
}
int mpas_b200_dycore_timer_reset(void* h) {
    Dycore* m = (Dycore*)h;
    for (auto& kv : m->timers) { kv.second.ms = 0; kv.second.calls = 0; }
    return 0;
}

}  // extern "C"
