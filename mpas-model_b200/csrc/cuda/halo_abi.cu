// Layer-1 halo exchange: the drop-in for the BODIES of mpas_dmpar_exch_halo_field2d_real / 3d_real
// (src/framework/mpas_dmpar.F:5386-5552, 5565-5748) for a Fortran host that keeps its own blocks, pools and exchange
// lists.  The host hands over, once, the send / recv lists it built with mpas_dmpar_get_exch_list
// (mpas_dmpar.F:1483-2038: per neighbour endPointID and halo layer, srcList / destList, 1-based) and the NCCL
// communicator id; afterwards mpas_b200_exch_halo_field2d / 3d take the same array the reference routine takes (a host
// address registered with mpas_b200_register_field, or a device pointer) plus the halo-layer selection of the optional
// haloLayersIn argument, and run  pack kernel -> ncclGroupStart; ncclRecv / ncclSend per neighbour; ncclGroupEnd ->
// unpack kernel  on the library stream, on the device mirror.  Lists whose endPoint is this rank are the reference's
// copyList (blocks of the same process, mpas_dmpar.F:5481-5504): a device copy.  One block per process (the reference's
// default config_number_of_blocks = 0).
#include <map>
#include <tuple>

#include "halo.h"

namespace mpasb200 {

int halo_exchange_messages(Nccl& nccl, bool have_comm, int my_rank, HaloPlan* p) {
    // self lists: sender segment -> receiver segment of the same plan
    bool remote = false;
    for (int r : p->send.nbr) remote = remote || r != my_rank;
    for (int r : p->recv.nbr) remote = remote || r != my_rank;
    if (remote) {
        if (!have_comm) return fail_msg("halo exchange: lists name other ranks but there is no communicator (mpas_b200_halo_comm_init)");
        if (nccl.group_start()) return fail_msg(std::string("ncclGroupStart: ") + nccl.last_error());
        for (size_t i = 0; i < p->recv.nbr.size(); ++i)
            if (p->recv.nbr[i] != my_rank && nccl.recv(p->recv.d_buf + p->recv.off[i], (size_t)p->recv.cnt[i], p->recv.nbr[i], stream()))
                return fail_msg(std::string("ncclRecv: ") + nccl.last_error());
        for (size_t i = 0; i < p->send.nbr.size(); ++i)
            if (p->send.nbr[i] != my_rank && nccl.send(p->send.d_buf + p->send.off[i], (size_t)p->send.cnt[i], p->send.nbr[i], stream()))
                return fail_msg(std::string("ncclSend: ") + nccl.last_error());
        if (nccl.group_end()) return fail_msg(std::string("ncclGroupEnd: ") + nccl.last_error());
        count_launch();
    }
    for (size_t i = 0; i < p->send.nbr.size(); ++i) {
        if (p->send.nbr[i] != my_rank) continue;
        size_t j = 0;
        for (; j < p->recv.nbr.size(); ++j)
            if (p->recv.nbr[j] == my_rank) break;
        if (j == p->recv.nbr.size() || p->recv.cnt[j] != p->send.cnt[i]) return fail_msg("halo exchange: unmatched self (copy) lists");
        cudaError_t e = cudaMemcpyAsync(p->recv.d_buf + p->recv.off[j], p->send.d_buf + p->send.off[i], (size_t)p->send.cnt[i] * sizeof(real),
                                        cudaMemcpyDeviceToDevice, stream());
        if (e != cudaSuccess) return fail("halo exchange copy", e);
        count_launch();
    }
    return 0;
}

namespace {
struct L1Halo {
    HaloBlock blk;
    Nccl nccl;
    bool have_comm = false;
    int rank = 0, nranks = 1;
    std::map<std::tuple<const void*, int, int, int>, HaloPlan*> plans;  // (device array, inner, kind, mask)
    long exchanges = 0;
};
L1Halo g_l1;

void l1_drop_plans() {
    for (auto& kv : g_l1.plans) halo_free_plan(kv.second);
    g_l1.plans.clear();
}

int l1_exchange(real* array, long long inner, int kind, int mask) {
    if (kind < 1 || kind > 3) return fail_msg("mpas_b200_exch_halo_field: kind must be 1 (cell), 2 (edge) or 3 (vertex)");
    if (inner < 1 || inner > 0x7fffffff) return fail_msg("mpas_b200_exch_halo_field: bad leading dimensions");
    real* d = X(array);
    const auto key = std::make_tuple((const void*)d, (int)inner, kind, mask);
    auto it = g_l1.plans.find(key);
    if (it == g_l1.plans.end()) {
        cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream(), &st) == cudaSuccess && st != cudaStreamCaptureStatusNone)
            return fail_msg("mpas_b200_exch_halo_field: first exchange of an array must be outside stream capture");
        HaloItem hi;
        hi.array = d; hi.loc = kind; hi.inner = (int)inner; hi.mask = mask;
        it = g_l1.plans.emplace(key, halo_build_plan(g_l1.blk, {hi})).first;
    }
    HaloPlan* p = it->second;
    ++g_l1.exchanges;
    int rc = halo_pack(p);
    if (!rc) rc = halo_exchange_messages(g_l1.nccl, g_l1.have_comm, g_l1.rank, p);
    if (!rc) rc = halo_unpack(p);
    return rc;
}
}  // namespace
}  // namespace mpasb200

using namespace mpasb200;

extern "C" {

// rank / size of the host's communicator and the NCCL id rank 0 obtained from mpas_b200_nccl_unique_id and broadcast
// (MPI_Bcast on the host side); nranks == 1 needs no id.
int mpas_b200_halo_comm_init(int rank, int nranks, const void* id128) {
    g_l1.rank = rank;
    g_l1.nranks = nranks;
    if (nranks > 1) {
        if (!id128) return fail_msg("mpas_b200_halo_comm_init: nranks > 1 needs the NCCL unique id");
        std::string err;
        if (!g_l1.nccl.load(err)) return fail_msg(err);
        if (g_l1.nccl.init_rank(id128, nranks, rank)) return fail_msg(std::string("ncclCommInitRank: ") + g_l1.nccl.last_error());
        g_l1.have_comm = true;
    }
    return 0;
}

// One call per (kind, side): all lists of that side, as mpas_dmpar_get_exch_list left them in
// field % sendList / recvList / copyList (mpas_dmpar_types.inc: endPointID, nList, srcList, destList; one list per halo
// layer in mpas_multihalo_exchange_list % halos(layer)).  which: 0 = send, 1 = recv.  srcList / destList are the nLists
// lists concatenated.  Replaces the lists registered before for that (kind, side) and drops the cached plans.
int mpas_b200_halo_register_lists(int kind, int which, int nLists, const int* endPoint, const int* haloLayer, const int* nList, const int* srcList,
                                  const int* destList) {
    if (kind < 1 || kind > 3 || which < 0 || which > 1 || nLists < 0) return fail_msg("mpas_b200_halo_register_lists: bad kind / which / nLists");
    std::vector<HaloList>& dst = which == 0 ? g_l1.blk.send[kind] : g_l1.blk.recv[kind];
    dst.clear();
    size_t cur = 0;
    for (int i = 0; i < nLists; ++i) {
        if (nList[i] < 0 || haloLayer[i] < 1 || haloLayer[i] > 30) return fail_msg("mpas_b200_halo_register_lists: bad list length or halo layer");
        HaloList l;
        l.endPoint = endPoint[i];
        l.layer = haloLayer[i];
        l.src.assign(srcList + cur, srcList + cur + nList[i]);
        l.dest.assign(destList + cur, destList + cur + nList[i]);
        for (int v : l.src) if (v < 1) return fail_msg("mpas_b200_halo_register_lists: list indices are 1-based");
        for (int v : l.dest) if (v < 1) return fail_msg("mpas_b200_halo_register_lists: list indices are 1-based");
        cur += (size_t)nList[i];
        dst.push_back(std::move(l));
    }
    cudaStreamSynchronize(stream());
    l1_drop_plans();
    return 0;
}

// mpas_dmpar_exch_halo_field2d_real(field, haloLayersIn), mpas_dmpar.F:5386: array(dim1, nElems+1), halo columns of the
// selected layers are overwritten.  layer_mask: bit (l-1) selects halo layer l, 0 = all layers (haloLayersIn absent).
int mpas_b200_exch_halo_field2d(mpas_real* array, int dim1, int kind, int layer_mask) {
    try {
        return l1_exchange(array, dim1, kind, layer_mask);
    } catch (const std::exception& e) { return fail_msg(e.what()); }
}
// mpas_dmpar_exch_halo_field3d_real, mpas_dmpar.F:5565: array(dim1, dim2, nElems+1)
int mpas_b200_exch_halo_field3d(mpas_real* array, int dim1, int dim2, int kind, int layer_mask) {
    try {
        return l1_exchange(array, (long long)dim1 * dim2, kind, layer_mask);
    } catch (const std::exception& e) { return fail_msg(e.what()); }
}
// 1-D decomposed fields (mpas_dmpar_exch_halo_field1d_real, mpas_dmpar.F:5253)
int mpas_b200_exch_halo_field1d(mpas_real* array, int kind, int layer_mask) {
    try {
        return l1_exchange(array, 1, kind, layer_mask);
    } catch (const std::exception& e) { return fail_msg(e.what()); }
}

long mpas_b200_halo_exchange_count(void) { return g_l1.exchanges; }

// The two exchanges inside atm_advance_scalars_mono_work (TI:3757 scalars_old all layers; TI:4098 scale_arr halo 1) as the
// callback mpas_b200_advance_scalars_mono_work takes: ctx = {nVertLevels, num_scalars} of the call (mpas_b200_mono_halo_ctx).
int mpas_b200_mono_halo_default(void* ctx, int which, mpas_real* array) {
    const int* c = (const int*)ctx;
    if (!c) return fail_msg("mpas_b200_mono_halo_default: ctx = {nVertLevels, num_scalars} required");
    const int L = c[0], ns = c[1];
    if (which == 0) return mpas_b200_exch_halo_field3d(array, ns, L, 1, 0);          // scalars(num_scalars, nVertLevels, nCells+1)
    if (which == 1) return mpas_b200_exch_halo_field3d(array, L, 2, 1, 1);           // scale_arr(nVertLevels, 2, nCells+1), (/1/)
    if (which == 2) return mpas_b200_exch_halo_field2d(array, 2 * ns * L, 1, 1);     // all-scalars scale factors, (/1/)
    return fail_msg("mpas_b200_mono_halo_default: unknown exchange");
}

int mpas_b200_halo_finalize(void) {
    cudaStreamSynchronize(stream());
    l1_drop_plans();
    for (int k = 0; k < 4; ++k) { g_l1.blk.send[k].clear(); g_l1.blk.recv[k].clear(); }
    g_l1.nccl.destroy();
    g_l1.have_comm = false;
    g_l1.nranks = 1;
    g_l1.rank = 0;
    return 0;
}

}  // extern "C"
