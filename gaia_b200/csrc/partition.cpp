// Slab partitioning of one tet mesh for the multi-GPU path - host code behind the C ABI
// (include/gaia_b200.h: gaia_b200_partition_*). The reference is single-device (M/VBD/VBDPhysics.h:212: one stream on one
// device), so this layer has no counterpart there; it decides WHO owns WHAT:
//   * owner of a vertex = its slab along the longest axis (equal vertex counts; vertices sharing a coordinate stay together);
//   * a partition holds every tet with at least one owned vertex, hence a one-ring GHOST layer of foreign vertices: owned
//     vertices see their complete stencil and the VBD update needs no force reduction;
//   * local vertices / tets keep their global order (ascending global id), so per-vertex incident-tet order, neighbour order
//     and with them every floating-point sum are the same as on one GPU;
//   * send[q] = owned vertices that are ghosts on rank q, recv[q] = ghosts owned by q - both ascending by global id, so
//     send[q] on this rank and recv[this] on rank q enumerate the same vertices in the same order.
// gaia_b200/multigpu.py binds these calls (bench, tests); gaia_b200/host/VBDPhysicsB200 uses them for `--gpus N`.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>

#include "../../include/gaia_b200.h"

struct gaia_partition {
    int32_t rank = 0, world = 1;
    std::vector<int32_t> localVerts;            // global ids, ascending
    std::vector<uint8_t> owned;                 // per local vertex
    std::vector<int64_t> localTets;             // global tet ids, ascending
    std::vector<int32_t> tets;                  // 4 * nLocalT, local vertex ids
    std::vector<int32_t> g2l;                   // global -> local or -1
    std::vector<uint8_t> boundary;              // per local vertex: sent to some peer
    struct Peer { int32_t rank; std::vector<int32_t> send, recv; };   // local vertex ids
    std::vector<Peer> peers;                    // ascending rank, only ranks with something to exchange
};

extern "C" {

int gaia_b200_partition_slab_owner(int32_t nV, const float* xyz, int32_t world, int32_t axis, int32_t* owner) {
    if (nV <= 0 || !xyz || !owner || world < 1) return GAIA_ERR_INVALID_ARG;
    if (axis < 0) {
        float lo[3], hi[3];
        for (int d = 0; d < 3; d++) lo[d] = hi[d] = xyz[d];
        for (int64_t v = 1; v < nV; v++) for (int d = 0; d < 3; d++) { lo[d] = std::min(lo[d], xyz[3 * v + d]); hi[d] = std::max(hi[d], xyz[3 * v + d]); }
        axis = 0;
        for (int d = 1; d < 3; d++) if (hi[d] - lo[d] > hi[axis] - lo[axis]) axis = d;
    }
    if (axis > 2) return GAIA_ERR_INVALID_ARG;
    std::vector<int32_t> order(nV);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return xyz[3 * (size_t)a + axis] < xyz[3 * (size_t)b + axis]; });
    std::vector<int64_t> bounds{0};
    for (int r = 1; r < world; r++) {
        int64_t cut = ((int64_t)nV * r) / world;
        // move the cut to the next change of coordinate so that a coordinate layer is never split
        while (cut < nV && cut > 0 && xyz[3 * (size_t)order[cut] + axis] == xyz[3 * (size_t)order[cut - 1] + axis]) cut++;
        bounds.push_back(cut);
    }
    bounds.push_back(nV);
    for (int r = 0; r < world; r++)
        for (int64_t i = bounds[r]; i < bounds[r + 1]; i++) owner[order[i]] = r;
    return GAIA_OK;
}

int gaia_b200_partition_create(int32_t nV, int64_t nT, const int32_t* tets, const int32_t* owner, int32_t rank, int32_t world, gaia_partition** out) {
    if (nV <= 0 || nT < 0 || (nT && !tets) || !owner || !out || rank < 0 || rank >= world) return GAIA_ERR_INVALID_ARG;
    gaia_partition* p = new gaia_partition();
    p->rank = rank; p->world = world;
    std::vector<uint8_t> used(nV, 0);
    for (int64_t t = 0; t < nT; t++) {
        bool mine = false;
        for (int k = 0; k < 4; k++) {
            const int32_t v = tets[4 * t + k];
            if (v < 0 || v >= nV) { delete p; return GAIA_ERR_INVALID_ARG; }
            mine |= owner[v] == rank;
        }
        if (!mine) continue;
        p->localTets.push_back(t);
        for (int k = 0; k < 4; k++) used[tets[4 * t + k]] = 1;
    }
    p->g2l.assign(nV, -1);
    for (int32_t v = 0; v < nV; v++) if (used[v]) { p->g2l[v] = (int32_t)p->localVerts.size(); p->localVerts.push_back(v); }
    const size_t nL = p->localVerts.size();
    p->owned.resize(nL);
    for (size_t i = 0; i < nL; i++) p->owned[i] = owner[p->localVerts[i]] == rank;
    p->tets.resize(4 * p->localTets.size());
    for (size_t i = 0; i < p->localTets.size(); i++)
        for (int k = 0; k < 4; k++) p->tets[4 * i + k] = p->g2l[tets[4 * p->localTets[i] + k]];
    // per peer: owned vertices that share a local tet with a vertex owned by the peer (= the peer's ghosts owned here), and
    // the local ghosts owned by the peer
    std::vector<std::vector<uint8_t>> sendMark(world);
    p->boundary.assign(nL, 0);
    for (size_t i = 0; i < p->localTets.size(); i++) {
        const int32_t* tv = &tets[4 * p->localTets[i]];
        for (int a = 0; a < 4; a++) {
            const int32_t q = owner[tv[a]];
            if (q == rank) continue;
            if (sendMark[q].empty()) sendMark[q].assign(nL, 0);
            for (int b = 0; b < 4; b++) if (owner[tv[b]] == rank) sendMark[q][p->g2l[tv[b]]] = 1;
        }
    }
    for (int q = 0; q < world; q++) {
        if (q == rank) continue;
        gaia_partition::Peer pe;
        pe.rank = q;
        if (!sendMark[q].empty()) for (size_t i = 0; i < nL; i++) if (sendMark[q][i]) { pe.send.push_back((int32_t)i); p->boundary[i] = 1; }
        for (size_t i = 0; i < nL; i++) if (!p->owned[i] && owner[p->localVerts[i]] == q) pe.recv.push_back((int32_t)i);
        if (!pe.send.empty() || !pe.recv.empty()) p->peers.push_back(std::move(pe));
    }
    *out = p;
    return GAIA_OK;
}

void gaia_b200_partition_destroy(gaia_partition* p) { delete p; }

int gaia_b200_partition_sizes(const gaia_partition* p, int32_t* nLocalV, int64_t* nLocalT, int32_t* nPeers) {
    if (!p) return GAIA_ERR_INVALID_ARG;
    if (nLocalV) *nLocalV = (int32_t)p->localVerts.size();
    if (nLocalT) *nLocalT = (int64_t)p->localTets.size();
    if (nPeers) *nPeers = (int32_t)p->peers.size();
    return GAIA_OK;
}
int gaia_b200_partition_local_verts(const gaia_partition* p, int32_t* globalIds, uint8_t* owned) {
    if (!p) return GAIA_ERR_INVALID_ARG;
    if (globalIds) std::copy(p->localVerts.begin(), p->localVerts.end(), globalIds);
    if (owned) std::copy(p->owned.begin(), p->owned.end(), owned);
    return GAIA_OK;
}
int gaia_b200_partition_local_tets(const gaia_partition* p, int32_t* tets, int64_t* globalTetIds) {
    if (!p) return GAIA_ERR_INVALID_ARG;
    if (tets) std::copy(p->tets.begin(), p->tets.end(), tets);
    if (globalTetIds) std::copy(p->localTets.begin(), p->localTets.end(), globalTetIds);
    return GAIA_OK;
}
int gaia_b200_partition_peer(const gaia_partition* p, int32_t i, int32_t* peerRank, int32_t* nSend, int32_t* nRecv) {
    if (!p || i < 0 || i >= (int32_t)p->peers.size()) return GAIA_ERR_INVALID_ARG;
    if (peerRank) *peerRank = p->peers[i].rank;
    if (nSend) *nSend = (int32_t)p->peers[i].send.size();
    if (nRecv) *nRecv = (int32_t)p->peers[i].recv.size();
    return GAIA_OK;
}
int gaia_b200_partition_peer_lists(const gaia_partition* p, int32_t i, int32_t* sendLocal, int32_t* recvLocal) {
    if (!p || i < 0 || i >= (int32_t)p->peers.size()) return GAIA_ERR_INVALID_ARG;
    if (sendLocal) std::copy(p->peers[i].send.begin(), p->peers[i].send.end(), sendLocal);
    if (recvLocal) std::copy(p->peers[i].recv.begin(), p->peers[i].recv.end(), recvLocal);
    return GAIA_OK;
}
// global ids -> local ids of the ones this partition holds (order kept); returns how many
int gaia_b200_partition_restrict_ids(const gaia_partition* p, int32_t n, const int32_t* globalIds, int32_t* localOut) {
    if (!p || n < 0 || (n && !globalIds)) return GAIA_ERR_INVALID_ARG;
    int32_t k = 0;
    for (int32_t i = 0; i < n; i++) {
        const int32_t g = globalIds[i];
        if (g < 0 || g >= (int32_t)p->g2l.size()) return GAIA_ERR_INVALID_ARG;
        if (p->g2l[g] >= 0) { if (localOut) localOut[k] = p->g2l[g]; k++; }
    }
    return k;
}
// global colouring -> local colouring, colour order kept (every partition sweeps the same colour together); inside a colour
// the vertices this rank SENDS sit behind the first 2/5 of the interior vertices: when their tiles come up the neighbours have
// long finished the previous colour (no stall on the epoch flag), and the halo stores still overlap the rest of the interior.
// outOffsets: nColors + 1; outVerts: sized for all local vertices.
int gaia_b200_partition_restrict_colors(const gaia_partition* p, int32_t nColors, const int32_t* offsets, const int32_t* verts,
                                        int32_t* outOffsets, int32_t* outVerts) {
    if (!p || nColors < 0 || !offsets || !outOffsets || !outVerts) return GAIA_ERR_INVALID_ARG;
    int32_t run = 0;
    std::vector<int32_t> inner, edge;
    for (int32_t c = 0; c < nColors; c++) {
        outOffsets[c] = run;
        inner.clear(); edge.clear();
        for (int32_t i = offsets[c]; i < offsets[c + 1]; i++) {
            const int32_t g = verts[i];
            if (g < 0 || g >= (int32_t)p->g2l.size()) return GAIA_ERR_INVALID_ARG;
            const int32_t l = p->g2l[g];
            if (l >= 0) (p->boundary[l] ? edge : inner).push_back(l);
        }
        const size_t k = (2 * inner.size()) / 5;
        for (size_t i = 0; i < k; i++) outVerts[run++] = inner[i];
        for (int32_t l : edge) outVerts[run++] = l;
        for (size_t i = k; i < inner.size(); i++) outVerts[run++] = inner[i];
    }
    outOffsets[nColors] = run;
    return GAIA_OK;
}

}  // extern "C"
