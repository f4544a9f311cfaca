// api_multi.inl -- slab-sharded rotation sweeps: merge the 3D components of contiguous frame slabs
// that were labelled on different GPUs.
//
// The reference labels a whole sweep on one host after the last frame
// (ConnectedComponents::find_3d_components, connected_components.cc:270-470).  Across a slab
// boundary a component can only continue through pixels with the SAME linear index in the last
// frame of slab r and the first frame of slab r+1 (cc:352-370), so only components that touch a
// slab's first or last plane ("boundary components") can change.  The exchange is therefore:
//   1. every rank sends its LAST plane {linear index, component} to rank r+1  (ncclSend/ncclRecv),
//      which intersects it with its own FIRST plane -> cross-slab equivalence pairs;
//   2. all ranks all-gather {pairs, boundary-component records} (ncclAllGather): a record is the
//      component's ffs_spot + its exact integer sums + second moments -- KBs;
//   3. every rank runs the same small union-find over the boundary components and rebuilds the
//      merged spots from the exact sums with the reference's arithmetic (double divide -> float
//      centre of mass, float peak-centroid distance, hpp:71-103,111-206).  A merged component
//      belongs to the slab that holds its canonical root = minimum (z, linear index) (cc:292-309,
//      396-399), so concatenating the ranks' lists in rank order IS the single-GPU order.
// NCCL is loaded at run time (dlopen of libnccl.so.2: inside a torch process that is the copy
// torch already loaded); a process that never creates an ffs_comm does not need it.
// ffs_sweep_merge_host is the same merge for slabs that live in one address space (several
// contexts in one process -- `spotfinder --devices`; also the CPU-testable form).
#include <dlfcn.h>
#include <nccl.h>

#include <chrono>
#include <map>
#include <unordered_map>

namespace ffsb {

struct BRec {  // boundary-component record (plain data, travels through ncclAllGather)
    uint32_t rank, comp;  // slab, and index of the component in that slab's canonical list
    ffs_spot spot;        // z still local to the slab
    uint64_t sums[4];     // sum I, (2x+1)I, (2y+1)I, (2z+1)I   (z local)
    uint64_t mom[4];      // bbox-relative sum I u^2, I v^2, I u v, I w^2
};
struct BPair {  // component comp_lo of slab rank_lo continues as comp_hi of slab rank_lo + 1
    uint32_t rank_lo, comp_lo, comp_hi, pad;
};
struct SlabMeta {
    uint64_t z0, n_frames, n_spots, n_first, n_last, n_brec, n_pairs, pad;
};
struct SpotEdit {  // what the merge does to one of a slab's own components
    uint32_t comp;
    bool drop;        // folded into a component of an earlier slab
    ffs_spot spot;    // otherwise: the merged component (z global), with its merged sums / moments
    uint64_t sums[4], mom[4];
};

typedef __int128 i128;

// Reflection3D::center_of_mass / peak_centroid_distance from exact sums (hpp:71-103,111-206),
// the same operations in the same order as acc_to_spot (cc.cuh).
static void spot_from_sums(ffs_spot& s, const uint64_t* sums) {
    s.sum_intensity = sums[0];
    if (sums[0] == 0) {
        s.com_x = s.com_y = s.com_z = s.peak_centroid_distance = NAN;
        return;
    }
    const double wt = (double)sums[0];
    s.com_x = (float)(((double)sums[1] * 0.5) / wt);
    s.com_y = (float)(((double)sums[2] * 0.5) / wt);
    s.com_z = (float)(((double)sums[3] * 0.5) / wt);
    volatile float dx = ((float)s.peak_x + 0.5f) - s.com_x;  // volatile: no contraction / excess precision
    volatile float dy = ((float)s.peak_y + 0.5f) - s.com_y;
    volatile float dz = ((float)s.peak_z + 0.5f) - s.com_z;
    volatile float xx = dx * dx, yy = dy * dy, zz = dz * dz;
    volatile float sxy = xx + yy;
    volatile float sum = sxy + zz;
    s.peak_centroid_distance = sqrtf(sum);
}

// slab-local z -> sweep z
static void globalise(ffs_spot& s, uint64_t* sums, uint64_t z0, uint64_t px) {
    s.z_min += (int32_t)z0;
    s.z_max += (int32_t)z0;
    s.peak_z += (int32_t)z0;
    s.label += z0 * px;
    sums[3] += 2 * z0 * sums[0];
    spot_from_sums(s, sums);
}

// bbox-relative second moments -> moments about the origin (exact), and back
static void moments_abs(const ffs_spot& s, const uint64_t* sums, const uint64_t* m, i128* out) {
    const i128 S0 = sums[0];
    const i128 Ix = ((i128)sums[1] - S0) / 2, Iy = ((i128)sums[2] - S0) / 2, Iz = ((i128)sums[3] - S0) / 2;
    const i128 a = s.x_min, b = s.y_min, c = s.z_min;
    const i128 Iu = Ix - a * S0, Iv = Iy - b * S0, Iw = Iz - c * S0;
    out[0] = (i128)m[0] + 2 * a * Iu + a * a * S0;
    out[1] = (i128)m[1] + 2 * b * Iv + b * b * S0;
    out[2] = (i128)m[2] + a * Iv + b * Iu + a * b * S0;
    out[3] = (i128)m[3] + 2 * c * Iw + c * c * S0;
}
static void moments_rel(const ffs_spot& s, const uint64_t* sums, const i128* absm, uint64_t* out) {
    const i128 S0 = sums[0];
    const i128 Ix = ((i128)sums[1] - S0) / 2, Iy = ((i128)sums[2] - S0) / 2, Iz = ((i128)sums[3] - S0) / 2;
    const i128 a = s.x_min, b = s.y_min, c = s.z_min;
    out[0] = (uint64_t)(absm[0] - 2 * a * Ix + a * a * S0);
    out[1] = (uint64_t)(absm[1] - 2 * b * Iy + b * b * S0);
    out[2] = (uint64_t)(absm[2] - a * Iy - b * Ix + a * b * S0);
    out[3] = (uint64_t)(absm[3] - 2 * c * Iz + c * c * S0);
}

// Cross-slab pairs: both planes are {index, component} lists in ascending linear index.
static void intersect_planes(const uint32_t* lo_last, size_t n_lo, const uint32_t* hi_first, size_t n_hi,
                             uint32_t rank_lo, std::vector<BPair>& pairs) {
    std::vector<uint64_t> seen;
    size_t i = 0, j = 0;
    while (i < n_lo && j < n_hi) {
        const uint32_t a = lo_last[2 * i], b = hi_first[2 * j];
        if (a < b) ++i;
        else if (b < a) ++j;
        else {
            seen.push_back((uint64_t)lo_last[2 * i + 1] << 32 | hi_first[2 * j + 1]);
            ++i;
            ++j;
        }
    }
    std::sort(seen.begin(), seen.end());
    seen.erase(std::unique(seen.begin(), seen.end()), seen.end());
    for (uint64_t k : seen) pairs.push_back(BPair{rank_lo, (uint32_t)(k >> 32), (uint32_t)k, 0});
}

// The boundary components of one slab: those listed in its first or last plane.
static void boundary_records(uint32_t rank, const ffs_spot* spots, const uint64_t* sums, const uint64_t* mom,
                             const uint32_t* first, size_t n_first, const uint32_t* last, size_t n_last,
                             std::vector<BRec>& out) {
    std::vector<uint32_t> comps;
    for (size_t i = 0; i < n_first; ++i) comps.push_back(first[2 * i + 1]);
    for (size_t i = 0; i < n_last; ++i) comps.push_back(last[2 * i + 1]);
    std::sort(comps.begin(), comps.end());
    comps.erase(std::unique(comps.begin(), comps.end()), comps.end());
    for (uint32_t c : comps) {
        BRec r{};
        r.rank = rank;
        r.comp = c;
        r.spot = spots[c];
        memcpy(r.sums, sums + 4 * (size_t)c, sizeof r.sums);
        if (mom) memcpy(r.mom, mom + 4 * (size_t)c, sizeof r.mom);
        out.push_back(r);
    }
}

// Step 3: the same on every rank.  recs / pairs hold every slab's records (any order).  Returns
// the edits for slab `me`.  Components not mentioned keep their (globalised) values.
static std::vector<SpotEdit> merge_core(const std::vector<SlabMeta>& meta, std::vector<BRec> recs,
                                        const std::vector<BPair>& pairs, uint32_t me, uint64_t px) {
    const size_t n = recs.size();
    std::map<std::pair<uint32_t, uint32_t>, size_t> at;  // (rank, comp) -> record
    for (size_t i = 0; i < n; ++i) at[{recs[i].rank, recs[i].comp}] = i;
    std::vector<size_t> parent(n);
    for (size_t i = 0; i < n; ++i) parent[i] = i;
    auto find = [&](size_t i) {
        while (parent[i] != i) {
            parent[i] = parent[parent[i]];
            i = parent[i];
        }
        return i;
    };
    // root = smallest (rank, comp) = smallest (z, linear index): slabs are ordered by z0 and a
    // slab's components are in canonical order
    auto key = [&](size_t i) { return std::make_pair(recs[i].rank, recs[i].comp); };
    for (const BPair& p : pairs) {
        auto a = at.find({p.rank_lo, p.comp_lo}), b = at.find({p.rank_lo + 1, p.comp_hi});
        if (a == at.end() || b == at.end()) continue;
        size_t ra = find(a->second), rb = find(b->second);
        if (ra == rb) continue;
        if (key(rb) < key(ra)) std::swap(ra, rb);
        parent[rb] = ra;
    }
    std::vector<std::vector<size_t>> members(n);
    for (size_t i = 0; i < n; ++i) members[find(i)].push_back(i);
    std::vector<SpotEdit> edits;
    for (size_t r = 0; r < n; ++r) {
        if (members[r].size() < 2) continue;
        bool mine = false;
        for (size_t i : members[r]) mine |= recs[i].rank == me;
        if (!mine) continue;
        // my non-root members are folded into the root's component (possibly another of mine:
        // two components of one slab can meet through a neighbouring slab)
        for (size_t i : members[r])
            if (i != r && recs[i].rank == me) {
                SpotEdit e{};
                e.comp = recs[i].comp;
                e.drop = true;
                edits.push_back(e);
            }
        if (recs[r].rank != me) continue;
        // I hold the root: build the merged component
        SpotEdit e{};
        e.comp = recs[r].comp;
        e.drop = false;
        i128 absm[4] = {0, 0, 0, 0};
        bool first = true;
        for (size_t i : members[r]) {
            BRec g = recs[i];
            globalise(g.spot, g.sums, meta[g.rank].z0, px);
            i128 am[4];
            moments_abs(g.spot, g.sums, g.mom, am);
            for (int k = 0; k < 4; ++k) absm[k] += am[k];
            if (first) {
                e.spot = g.spot;
                memcpy(e.sums, g.sums, sizeof e.sums);
                first = false;
                continue;
            }
            ffs_spot& a = e.spot;
            const ffs_spot& b = g.spot;
            a.x_min = std::min(a.x_min, b.x_min);
            a.y_min = std::min(a.y_min, b.y_min);
            a.z_min = std::min(a.z_min, b.z_min);
            a.x_max = std::max(a.x_max, b.x_max);
            a.y_max = std::max(a.y_max, b.y_max);
            a.z_max = std::max(a.z_max, b.z_max);
            a.num_pixels += b.num_pixels;
            a.label = std::min(a.label, b.label);
            // peak: brightest, ties -> smallest (z, y, x)   (connected_components.cc:143-157)
            const auto ka = std::make_tuple(~a.peak_intensity, a.peak_z, a.peak_y, a.peak_x);
            const auto kb = std::make_tuple(~b.peak_intensity, b.peak_z, b.peak_y, b.peak_x);
            if (kb < ka) {
                a.peak_intensity = b.peak_intensity;
                a.peak_x = b.peak_x;
                a.peak_y = b.peak_y;
                a.peak_z = b.peak_z;
            }
            for (int k = 0; k < 4; ++k) e.sums[k] += g.sums[k];
        }
        spot_from_sums(e.spot, e.sums);
        moments_rel(e.spot, e.sums, absm, e.mom);
        edits.push_back(e);
    }
    std::sort(edits.begin(), edits.end(), [](const SpotEdit& a, const SpotEdit& b) { return a.comp < b.comp; });
    return edits;
}

// Apply the edits to one slab's lists: globalise everything, replace merged roots, drop folded
// members.  Order is preserved, which is the canonical order.
static void apply_edits(const std::vector<SpotEdit>& edits, uint64_t z0, uint64_t px, const ffs_spot* spots,
                        const uint64_t* sums, const uint64_t* mom, size_t n, std::vector<ffs_spot>& out_spots,
                        std::vector<uint64_t>& out_sums, std::vector<uint64_t>& out_mom) {
    size_t e = 0;
    for (size_t c = 0; c < n; ++c) {
        while (e < edits.size() && edits[e].comp < c) ++e;
        if (e < edits.size() && edits[e].comp == c) {
            if (edits[e].drop) continue;
            out_spots.push_back(edits[e].spot);
            out_sums.insert(out_sums.end(), edits[e].sums, edits[e].sums + 4);
            out_mom.insert(out_mom.end(), edits[e].mom, edits[e].mom + 4);
            continue;
        }
        ffs_spot s = spots[c];
        uint64_t sm[4];
        memcpy(sm, sums + 4 * c, sizeof sm);
        globalise(s, sm, z0, px);
        out_spots.push_back(s);
        out_sums.insert(out_sums.end(), sm, sm + 4);
        if (mom) out_mom.insert(out_mom.end(), mom + 4 * c, mom + 4 * c + 4);
        else out_mom.insert(out_mom.end(), 4, 0ull);
    }
}

// ---- NCCL at run time ----------------------------------------------------------------------
struct NcclApi {
    void* handle = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    std::string err;
};

static NcclApi* nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) {
            api.err = std::string("libnccl.so.2 not found: ") + (dlerror() ? dlerror() : "");
            return;
        }
#define FFS_NCCL_SYM(f)                                                         \
    api.f = reinterpret_cast<decltype(api.f)>(dlsym(api.handle, "nccl" #f));    \
    if (!api.f) api.err = "libnccl: missing symbol nccl" #f;
        FFS_NCCL_SYM(GetUniqueId)
        FFS_NCCL_SYM(CommInitRank)
        FFS_NCCL_SYM(CommDestroy)
        FFS_NCCL_SYM(AllGather)
        FFS_NCCL_SYM(Send)
        FFS_NCCL_SYM(Recv)
        FFS_NCCL_SYM(GroupStart)
        FFS_NCCL_SYM(GroupEnd)
        FFS_NCCL_SYM(GetErrorString)
#undef FFS_NCCL_SYM
    });
    return api.err.empty() ? &api : nullptr;
}

}  // namespace ffsb

struct ffs_comm {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1, device = 0;
    cudaStream_t stream = nullptr;
    uint8_t *d_send = nullptr, *d_recv = nullptr;
    size_t cap_send = 0, cap_recv = 0;
    std::string err;
};

static thread_local std::string g_comm_error;

#define COMM_CUDA(cm, call)                                                         \
    do {                                                                            \
        cudaError_t e__ = (call);                                                   \
        if (e__ != cudaSuccess) {                                                   \
            (cm)->err = std::string(#call) + ": " + cudaGetErrorString(e__);        \
            return FFS_ERR_CUDA;                                                    \
        }                                                                           \
    } while (0)
#define COMM_NCCL(cm, call)                                                         \
    do {                                                                            \
        ncclResult_t r__ = (call);                                                  \
        if (r__ != ncclSuccess) {                                                   \
            (cm)->err = std::string(#call) + ": " + nccl_api()->GetErrorString(r__); \
            return FFS_ERR_CUDA;                                                    \
        }                                                                           \
    } while (0)

static int comm_allgather(ffs_comm* cm, const void* send, size_t bytes, void* recv_all);
static int comm_shift_up(ffs_comm* cm, const void* send, size_t n_send, void* recv, size_t n_recv);
extern "C" void ffs_comm_destroy(ffs_comm* cm);

extern "C" int ffs_comm_unique_id(uint8_t* id128) {
    NcclApi* n = nccl_api();
    if (!n || !id128) return FFS_ERR_ARG;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if (n->GetUniqueId(&id) != ncclSuccess) return FFS_ERR_CUDA;
    memcpy(id128, &id, 128);
    return FFS_OK;
}

extern "C" const char* ffs_comm_last_error(const ffs_comm* cm) { return cm ? cm->err.c_str() : g_comm_error.c_str(); }

extern "C" int ffs_comm_create(const uint8_t* id128, int rank, int world, int device, ffs_comm** out) {
    if (!id128 || !out || world < 1 || rank < 0 || rank >= world) return FFS_ERR_ARG;
    *out = nullptr;
    NcclApi* n = nccl_api();
    if (!n) {
        g_comm_error = "NCCL is not available in this process";
        return FFS_ERR_NO_DEVICE;
    }
    ffs_comm* cm = new (std::nothrow) ffs_comm();
    if (!cm) return FFS_ERR_NOMEM;
    cm->rank = rank;
    cm->world = world;
    cm->device = device;
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&cm->stream, cudaStreamNonBlocking);
    ncclResult_t r = e == cudaSuccess ? n->CommInitRank(&cm->comm, world, id, rank) : ncclSuccess;
    if (e != cudaSuccess || r != ncclSuccess) {
        g_comm_error = e != cudaSuccess ? cudaGetErrorString(e) : n->GetErrorString(r);
        delete cm;
        return FFS_ERR_CUDA;
    }
    // NCCL sets its channels and peer connections up lazily, at the first collective / point-to-point
    // call (1.9 s for 8 ranks, measured): do that here, not inside the first ffs_sweep_merge
    {
        std::vector<uint64_t> all(world);
        uint64_t mine = (uint64_t)rank, prev = 0;
        int rc = comm_allgather(cm, &mine, sizeof mine, all.data());
        if (rc == FFS_OK) rc = comm_shift_up(cm, &mine, sizeof mine, &prev, sizeof prev);
        if (rc != FFS_OK) {
            g_comm_error = cm->err;
            ffs_comm_destroy(cm);
            return rc;
        }
    }
    *out = cm;
    return FFS_OK;
}

extern "C" void ffs_comm_destroy(ffs_comm* cm) {
    if (!cm) return;
    cudaSetDevice(cm->device);
    if (cm->comm && nccl_api()) nccl_api()->CommDestroy(cm->comm);
    cudaFree(cm->d_send);
    cudaFree(cm->d_recv);
    if (cm->stream) cudaStreamDestroy(cm->stream);
    delete cm;
}

extern "C" int ffs_comm_rank(const ffs_comm* cm) { return cm ? cm->rank : -1; }
extern "C" int ffs_comm_world(const ffs_comm* cm) { return cm ? cm->world : 0; }

static int comm_reserve(ffs_comm* cm, size_t send, size_t recv) {
    if (send > cm->cap_send) {
        cudaFree(cm->d_send);
        cm->cap_send = std::max<size_t>(send, 1 << 16);
        COMM_CUDA(cm, cudaMalloc(&cm->d_send, cm->cap_send));
    }
    if (recv > cm->cap_recv) {
        cudaFree(cm->d_recv);
        cm->cap_recv = std::max<size_t>(recv, 1 << 16);
        COMM_CUDA(cm, cudaMalloc(&cm->d_recv, cm->cap_recv));
    }
    return FFS_OK;
}

// every rank contributes `bytes` bytes; recv_all receives world * bytes, in rank order
static int comm_allgather(ffs_comm* cm, const void* send, size_t bytes, void* recv_all) {
    int rc = comm_reserve(cm, bytes, bytes * cm->world);
    if (rc) return rc;
    COMM_CUDA(cm, cudaMemcpyAsync(cm->d_send, send, bytes, cudaMemcpyHostToDevice, cm->stream));
    COMM_NCCL(cm, nccl_api()->AllGather(cm->d_send, cm->d_recv, bytes, ncclUint8, cm->comm, cm->stream));
    COMM_CUDA(cm, cudaMemcpyAsync(recv_all, cm->d_recv, bytes * cm->world, cudaMemcpyDeviceToHost, cm->stream));
    COMM_CUDA(cm, cudaStreamSynchronize(cm->stream));
    return FFS_OK;
}

// rank r sends n_send bytes to r+1 and receives n_recv bytes from r-1 (grouped point-to-point)
static int comm_shift_up(ffs_comm* cm, const void* send, size_t n_send, void* recv, size_t n_recv) {
    int rc = comm_reserve(cm, n_send, n_recv);
    if (rc) return rc;
    const bool do_send = cm->rank + 1 < cm->world && n_send > 0, do_recv = cm->rank > 0 && n_recv > 0;
    if (do_send) COMM_CUDA(cm, cudaMemcpyAsync(cm->d_send, send, n_send, cudaMemcpyHostToDevice, cm->stream));
    NcclApi* n = nccl_api();
    COMM_NCCL(cm, n->GroupStart());
    if (do_send) COMM_NCCL(cm, n->Send(cm->d_send, n_send, ncclUint8, cm->rank + 1, cm->comm, cm->stream));
    if (do_recv) COMM_NCCL(cm, n->Recv(cm->d_recv, n_recv, ncclUint8, cm->rank - 1, cm->comm, cm->stream));
    COMM_NCCL(cm, n->GroupEnd());
    if (do_recv) COMM_CUDA(cm, cudaMemcpyAsync(recv, cm->d_recv, n_recv, cudaMemcpyDeviceToHost, cm->stream));
    COMM_CUDA(cm, cudaStreamSynchronize(cm->stream));
    return FFS_OK;
}

// Distributed merge: call on every rank after ffs_sweep_finish.
extern "C" int ffs_sweep_merge(ffs_ctx* c, ffs_comm* cm, uint64_t z0, uint64_t n_frames, const ffs_spot* spots,
                               size_t n_spots, ffs_spot** out_spots, size_t* n_out, double* exchange_seconds) {
    if (!c || !cm || !out_spots || !n_out || (n_spots && !spots)) return FFS_ERR_ARG;
    if (c->sweep_sums.size() != 4 * n_spots || c->sweep_moments.size() != 4 * n_spots) return FFS_ERR_STATE;
    *out_spots = nullptr;
    *n_out = 0;
    const auto t0 = std::chrono::steady_clock::now();
    if (cudaSetDevice(cm->device) != cudaSuccess) return FFS_ERR_CUDA;
    const uint64_t px = (uint64_t)c->p.width * c->p.height;
    const uint32_t me = (uint32_t)cm->rank, W = (uint32_t)cm->world;
    const auto& first = c->sweep_plane[0];
    const auto& last = c->sweep_plane[1];
    int rc;
#define MERGE_RC(call)          \
    if ((rc = (call)) != 0) {   \
        c->err = cm->err;       \
        return rc;              \
    }
    // ---- sizes ---------------------------------------------------------------------------------
    std::vector<SlabMeta> meta(W);
    SlabMeta mine{z0, n_frames, n_spots, first.size() / 2, last.size() / 2, 0, 0, 0};
    MERGE_RC(comm_allgather(cm, &mine, sizeof mine, meta.data()));
    // ---- 1. last plane -> next rank; intersect with my first plane --------------------------------
    std::vector<uint32_t> prev_last(me > 0 ? 2 * meta[me - 1].n_last : 0);
    MERGE_RC(comm_shift_up(cm, last.data(), last.size() * 4, prev_last.data(), prev_last.size() * 4));
    std::vector<BPair> my_pairs;
    if (me > 0 && meta[me - 1].z0 + meta[me - 1].n_frames == z0)  // really consecutive frames
        intersect_planes(prev_last.data(), prev_last.size() / 2, first.data(), first.size() / 2, me - 1, my_pairs);
    std::vector<BRec> my_recs;
    boundary_records(me, spots, c->sweep_sums.data(), c->sweep_moments.data(), first.data(), first.size() / 2,
                     last.data(), last.size() / 2, my_recs);
    // ---- 2. all-gather records and pairs (padded to the largest contribution) --------------------
    uint64_t cnt[2] = {my_recs.size(), my_pairs.size()};
    std::vector<uint64_t> cnts(2 * W);
    MERGE_RC(comm_allgather(cm, cnt, sizeof cnt, cnts.data()));
    size_t max_bytes = 16;
    for (uint32_t r = 0; r < W; ++r)
        max_bytes = std::max<size_t>(max_bytes, cnts[2 * r] * sizeof(BRec) + cnts[2 * r + 1] * sizeof(BPair));
    std::vector<uint8_t> sendbuf(max_bytes, 0), recvbuf(max_bytes * W);
    memcpy(sendbuf.data(), my_recs.data(), my_recs.size() * sizeof(BRec));
    memcpy(sendbuf.data() + my_recs.size() * sizeof(BRec), my_pairs.data(), my_pairs.size() * sizeof(BPair));
    MERGE_RC(comm_allgather(cm, sendbuf.data(), max_bytes, recvbuf.data()));
    std::vector<BRec> recs;
    std::vector<BPair> pairs;
    for (uint32_t r = 0; r < W; ++r) {
        const uint8_t* p = recvbuf.data() + (size_t)r * max_bytes;
        const BRec* br = reinterpret_cast<const BRec*>(p);
        recs.insert(recs.end(), br, br + cnts[2 * r]);
        const BPair* bp = reinterpret_cast<const BPair*>(p + cnts[2 * r] * sizeof(BRec));
        pairs.insert(pairs.end(), bp, bp + cnts[2 * r + 1]);
    }
#undef MERGE_RC
    if (exchange_seconds) *exchange_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    // ---- 3. merge --------------------------------------------------------------------------------
    const std::vector<SpotEdit> edits = merge_core(meta, recs, pairs, me, px);
    std::vector<ffs_spot> os;
    std::vector<uint64_t> osum, omom;
    os.reserve(n_spots);
    apply_edits(edits, z0, px, spots, c->sweep_sums.data(), c->sweep_moments.data(), n_spots, os, osum, omom);
    c->sweep_sums = std::move(osum);
    c->sweep_moments = std::move(omom);
    if (!os.empty()) {
        *out_spots = (ffs_spot*)malloc(os.size() * sizeof(ffs_spot));
        if (!*out_spots) return FFS_ERR_NOMEM;
        memcpy(*out_spots, os.data(), os.size() * sizeof(ffs_spot));
    }
    *n_out = os.size();
    return FFS_OK;
}

// Every rank's final list, concatenated in rank order (= canonical order), on every rank.
extern "C" int ffs_sweep_gather(ffs_comm* cm, const ffs_spot* mine, size_t n, ffs_spot** all, size_t* n_all) {
    if (!cm || !all || !n_all || (n && !mine)) return FFS_ERR_ARG;
    *all = nullptr;
    *n_all = 0;
    if (cudaSetDevice(cm->device) != cudaSuccess) return FFS_ERR_CUDA;
    const uint32_t W = (uint32_t)cm->world;
    uint64_t cnt = n;
    std::vector<uint64_t> cnts(W);
    int rc = comm_allgather(cm, &cnt, sizeof cnt, cnts.data());
    if (rc) return rc;
    size_t mx = 1, total = 0;
    for (uint64_t v : cnts) {
        mx = std::max<size_t>(mx, v);
        total += v;
    }
    std::vector<ffs_spot> sendbuf(mx), recvbuf(mx * W);
    if (n) memcpy(sendbuf.data(), mine, n * sizeof(ffs_spot));
    rc = comm_allgather(cm, sendbuf.data(), mx * sizeof(ffs_spot), recvbuf.data());
    if (rc) return rc;
    if (total == 0) return FFS_OK;
    ffs_spot* out = (ffs_spot*)malloc(total * sizeof(ffs_spot));
    if (!out) return FFS_ERR_NOMEM;
    size_t k = 0;
    for (uint32_t r = 0; r < W; ++r) {
        memcpy(out + k, recvbuf.data() + (size_t)r * mx, cnts[r] * sizeof(ffs_spot));
        k += cnts[r];
    }
    *all = out;
    *n_all = total;
    return FFS_OK;
}

// The same merge for slabs held in ONE address space (plain host arrays: needs no GPU).
extern "C" int ffs_sweep_merge_host(int n_slabs, uint32_t width, uint32_t height, const uint64_t* z0,
                                    const uint64_t* n_frames, const ffs_spot* const* spots, const size_t* n_spots,
                                    const uint64_t* const* sums, const uint64_t* const* moments,
                                    const uint32_t* const* first, const size_t* n_first,
                                    const uint32_t* const* last, const size_t* n_last, ffs_spot** out_spots,
                                    uint64_t** out_sums, uint64_t** out_moments, size_t* n_out) {
    if (n_slabs < 1 || !z0 || !n_frames || !spots || !n_spots || !sums || !first || !n_first || !last || !n_last ||
        !out_spots || !n_out)
        return FFS_ERR_ARG;
    *out_spots = nullptr;
    *n_out = 0;
    if (out_sums) *out_sums = nullptr;
    if (out_moments) *out_moments = nullptr;
    const uint64_t px = (uint64_t)width * height;
    std::vector<int> order(n_slabs);
    for (int i = 0; i < n_slabs; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int a, int b) { return z0[a] < z0[b]; });
    std::vector<SlabMeta> meta(n_slabs);
    std::vector<BRec> recs;
    std::vector<BPair> pairs;
    for (int r = 0; r < n_slabs; ++r) {
        const int s = order[r];
        meta[r] = SlabMeta{z0[s], n_frames[s], n_spots[s], n_first[s], n_last[s], 0, 0, 0};
        boundary_records((uint32_t)r, spots[s], sums[s], moments ? moments[s] : nullptr, first[s], n_first[s],
                         last[s], n_last[s], recs);
        if (r > 0 && meta[r - 1].z0 + meta[r - 1].n_frames == meta[r].z0) {
            const int q = order[r - 1];
            intersect_planes(last[q], n_last[q], first[s], n_first[s], (uint32_t)(r - 1), pairs);
        }
    }
    std::vector<ffs_spot> os;
    std::vector<uint64_t> osum, omom;
    for (int r = 0; r < n_slabs; ++r) {
        const int s = order[r];
        const std::vector<SpotEdit> edits = merge_core(meta, recs, pairs, (uint32_t)r, px);
        apply_edits(edits, meta[r].z0, px, spots[s], sums[s], moments ? moments[s] : nullptr, n_spots[s], os, osum,
                    omom);
    }
    if (os.empty()) return FFS_OK;
    *out_spots = (ffs_spot*)malloc(os.size() * sizeof(ffs_spot));
    if (!*out_spots) return FFS_ERR_NOMEM;
    memcpy(*out_spots, os.data(), os.size() * sizeof(ffs_spot));
    if (out_sums) {
        *out_sums = (uint64_t*)malloc(osum.size() * sizeof(uint64_t));
        if (*out_sums) memcpy(*out_sums, osum.data(), osum.size() * sizeof(uint64_t));
    }
    if (out_moments) {
        *out_moments = (uint64_t*)malloc(omom.size() * sizeof(uint64_t));
        if (*out_moments) memcpy(*out_moments, omom.data(), omom.size() * sizeof(uint64_t));
    }
    *n_out = os.size();
    return FFS_OK;
}
