#include "VBDPhysicsB200MultiGPU.h"

#include <algorithm>
#include <chrono>
#include <cstring>
#include <functional>
#include <iostream>
#include <stdexcept>

#include "host_util.h"

namespace gaia_b200 {

VBDPhysicsB200MultiGPU::~VBDPhysicsB200MultiGPU() {
    for (auto& pt : parts_) {
        if (pt.h) gaia_b200_destroy(pt.h);
        if (pt.p) gaia_b200_partition_destroy(pt.p);
    }
}

void VBDPhysicsB200MultiGPU::forAllParts(const std::function<void(int)>& fn) {
    std::vector<std::thread> th;
    std::vector<std::string> err(parts_.size());
    for (int r = 0; r < (int)parts_.size(); r++)
        th.emplace_back([&, r] { try { fn(r); } catch (const std::exception& e) { err[r] = e.what(); } });
    for (auto& t : th) t.join();
    for (size_t r = 0; r < err.size(); r++) if (!err[r].empty()) throw std::runtime_error("GPU " + std::to_string(r) + ": " + err[r]);
}

void VBDPhysicsB200MultiGPU::initialize() {
    if (models_.size() != 1) throw std::runtime_error("--gpus N partitions ONE tet mesh into slabs (many-object scenes are sharded by object: gaia_b200/multigpu.py)");
    if (meshCache_.empty()) loadMeshes(nullptr, nullptr);
    const ModelInput& m = models_[0];
    const HostMeshData& hm = meshCache_[m.meshCacheId];
    const int32_t nV = (int32_t)(hm.verts.size() / 3);
    const int64_t nT = (int64_t)(hm.tets.size() / 4);
    // global colour order: size-descending, ties keep file order (what the single-GPU path does inside add_tet_mesh)
    std::vector<std::vector<int32_t>> colours = colourings_.at(m.verticesColoringCategoriesPath);
    std::stable_sort(colours.begin(), colours.end(), [](const std::vector<int32_t>& a, const std::vector<int32_t>& b) { return a.size() > b.size(); });
    std::vector<int32_t> cOffs(1, 0), cFlat;
    for (auto& c : colours) { cFlat.insert(cFlat.end(), c.begin(), c.end()); cOffs.push_back((int32_t)cFlat.size()); }
    std::vector<int32_t> owner(nV);
    check(gaia_b200_partition_slab_owner(nV, hm.verts.data(), nGpus_, -1, owner.data()), "gaia_b200_partition_slab_owner");
    parts_.assign(nGpus_, Part());
    if (replicated_) {
        if (nGpus_ > 8) throw std::runtime_error("--layout replicated: at most 8 GPUs (every GPU is every other GPU's peer)");
        initializeReplicated(owner, cOffs, cFlat, (int)colours.size());
        global_.assign(3 * (size_t)nV, 0.f);
        frameId_ = 0;
        return;
    }
    forAllParts([&](int r) {
        Part& pt = parts_[r];
        pt.device = r;
        if (gaia_b200_partition_create(nV, nT, hm.tets.data(), owner.data(), r, nGpus_, &pt.p) < 0) throw std::runtime_error("gaia_b200_partition_create failed");
        int32_t nLV = 0; int64_t nLT = 0;
        gaia_b200_partition_sizes(pt.p, &nLV, &nLT, nullptr);
        pt.localVerts.resize(nLV); pt.owned.resize(nLV);
        gaia_b200_partition_local_verts(pt.p, pt.localVerts.data(), pt.owned.data());
        std::vector<int32_t> tets(4 * (size_t)nLT);
        gaia_b200_partition_local_tets(pt.p, tets.data(), nullptr);
        std::vector<float> xyz(3 * (size_t)nLV);
        for (int32_t i = 0; i < nLV; i++) std::memcpy(&xyz[3 * (size_t)i], &hm.verts[3 * (size_t)pt.localVerts[i]], 12);
        std::vector<int32_t> lo(colours.size() + 1), lv(std::max<int32_t>(nLV, 1));
        if (gaia_b200_partition_restrict_colors(pt.p, (int32_t)colours.size(), cOffs.data(), cFlat.data(), lo.data(), lv.data()) < 0)
            throw std::runtime_error("gaia_b200_partition_restrict_colors failed");
        std::vector<int32_t> fixed(std::max<size_t>(m.fixedPoints.size(), 1));
        const int nFixed = gaia_b200_partition_restrict_ids(pt.p, (int32_t)m.fixedPoints.size(), m.fixedPoints.data(), fixed.data());
        auto ck = [&](int rc, const char* what) { if (rc < 0) throw std::runtime_error(std::string(what) + ": " + gaia_b200_last_error(pt.h)); };
        ck(gaia_b200_create(&phys_, &col_, pt.device, &pt.h), "gaia_b200_create");
        if (phys_.useBowlGround) ck(gaia_b200_set_bowl_ground(pt.h, bowlCenter_, bowlRadius_, bowlCap_ ? 1 : 0), "gaia_b200_set_bowl_ground");
        ck(gaia_b200_set_partition_mode(pt.h, r, nGpus_), "gaia_b200_set_partition_mode");
        ck(gaia_b200_add_tet_mesh(pt.h, nLV, (int32_t)nLT, xyz.data(), tets.data(), &m.obj, (int32_t)colours.size(), lo.data(), lv.data(), nFixed,
                                  nFixed ? fixed.data() : nullptr), "gaia_b200_add_tet_mesh");
        std::vector<int32_t> ghosts;
        for (int32_t i = 0; i < nLV; i++) if (!pt.owned[i]) ghosts.push_back(i);
        ck(gaia_b200_set_ghosts(pt.h, 0, ghosts.data(), (int32_t)ghosts.size()), "gaia_b200_set_ghosts");
        for (auto& d : deformers_)
            for (size_t i = 0; i < d.selectedMeshes.size(); i++) {
                std::vector<int32_t> g(d.selectedVertices[i].begin(), d.selectedVertices[i].end()), l(std::max<size_t>(g.size(), 1));
                const int n = gaia_b200_partition_restrict_ids(pt.p, (int32_t)g.size(), g.data(), l.data());
                if (n <= 0) continue;
                if (d.type == 0) ck(gaia_b200_add_rotator(pt.h, 0, l.data(), n, d.center, d.axis, d.angularVelocity, d.endTime), "gaia_b200_add_rotator");
                else ck(gaia_b200_add_translator(pt.h, 0, l.data(), n, d.speed, d.endTime), "gaia_b200_add_translator");
            }
        ck(gaia_b200_finalize(pt.h), "gaia_b200_finalize");
    });
    // halo set-up (every partition is finalized): where my boundary vertices live on each neighbour = the device slots of the
    // neighbour's ghosts that I own, in the same (ascending global id) order as my send list
    std::vector<std::array<unsigned char, 16>> ptrs(nGpus_);
    for (int r = 0; r < nGpus_; r++) if (gaia_b200_halo_local_pointers(parts_[r].h, ptrs[r].data()) < 0) throw std::runtime_error(gaia_b200_last_error(parts_[r].h));
    for (int r = 0; r < nGpus_; r++) {
        Part& pt = parts_[r];
        int32_t nPeers = 0;
        gaia_b200_partition_sizes(pt.p, nullptr, nullptr, &nPeers);
        for (int i = 0; i < nPeers; i++) {
            int32_t q = 0, nSend = 0, nRecv = 0;
            gaia_b200_partition_peer(pt.p, i, &q, &nSend, &nRecv);
            if (!nSend) continue;
            std::vector<int32_t> send(nSend);
            gaia_b200_partition_peer_lists(pt.p, i, send.data(), nullptr);
            // the neighbour's recv list for me
            Part& pq = parts_[q];
            int32_t nPq = 0;
            gaia_b200_partition_sizes(pq.p, nullptr, nullptr, &nPq);
            std::vector<int32_t> remote;
            for (int j = 0; j < nPq; j++) {
                int32_t qq = 0, ns = 0, nr = 0;
                gaia_b200_partition_peer(pq.p, j, &qq, &ns, &nr);
                if (qq != r) continue;
                std::vector<int32_t> recv(std::max(nr, 1));
                gaia_b200_partition_peer_lists(pq.p, j, nullptr, recv.data());
                remote.resize(nr);
                if (gaia_b200_get_sorted_ids(pq.h, 0, recv.data(), nr, remote.data()) < 0) throw std::runtime_error(gaia_b200_last_error(pq.h));
            }
            if ((int32_t)remote.size() != nSend) throw std::runtime_error("halo tables of ranks " + std::to_string(r) + " and " + std::to_string(q) + " disagree");
            std::vector<int32_t> mv(2 * (size_t)nSend, 0);
            for (int k = 0; k < nSend; k++) mv[2 * k + 1] = send[k];
            if (gaia_b200_halo_add_peer_pointers(pt.h, q, pq.device, ptrs[q].data(), nSend, mv.data(), remote.data()) < 0)
                throw std::runtime_error(gaia_b200_last_error(pt.h));
        }
    }
    for (auto& pt : parts_) if (gaia_b200_halo_finalize(pt.h) < 0) throw std::runtime_error(gaia_b200_last_error(pt.h));
    global_.assign(3 * (size_t)nV, 0.f);
    frameId_ = 0;
}

// Replicated layout: full mesh on every GPU (global ids = local ids, global colour order), foreign vertices as ghosts, peers =
// all other GPUs, send list = my owned vertices, remote slot = the peer's slot of the same vertex.
void VBDPhysicsB200MultiGPU::initializeReplicated(const std::vector<int32_t>& owner, const std::vector<int32_t>& cOffs, const std::vector<int32_t>& cFlat,
                                                  int nColours) {
    const ModelInput& m = models_[0];
    const HostMeshData& hm = meshCache_[m.meshCacheId];
    const int32_t nV = (int32_t)(hm.verts.size() / 3);
    const int64_t nT = (int64_t)(hm.tets.size() / 4);
    forAllParts([&](int r) {
        Part& pt = parts_[r];
        pt.device = r;
        pt.localVerts.resize(nV); pt.owned.resize(nV);
        std::vector<int32_t> ghosts;
        for (int32_t v = 0; v < nV; v++) { pt.localVerts[v] = v; pt.owned[v] = owner[v] == r; if (owner[v] != r) ghosts.push_back(v); }
        auto ck = [&](int rc, const char* what) { if (rc < 0) throw std::runtime_error(std::string(what) + ": " + gaia_b200_last_error(pt.h)); };
        ck(gaia_b200_create(&phys_, &col_, pt.device, &pt.h), "gaia_b200_create");
        if (phys_.useBowlGround) ck(gaia_b200_set_bowl_ground(pt.h, bowlCenter_, bowlRadius_, bowlCap_ ? 1 : 0), "gaia_b200_set_bowl_ground");
        ck(gaia_b200_set_partition_mode(pt.h, r, nGpus_), "gaia_b200_set_partition_mode");
        ck(gaia_b200_add_tet_mesh(pt.h, nV, (int32_t)nT, hm.verts.data(), hm.tets.data(), &m.obj, nColours, cOffs.data(), cFlat.data(),
                                  (int32_t)m.fixedPoints.size(), m.fixedPoints.empty() ? nullptr : m.fixedPoints.data()), "gaia_b200_add_tet_mesh");
        ck(gaia_b200_set_ghosts(pt.h, 0, ghosts.data(), (int32_t)ghosts.size()), "gaia_b200_set_ghosts");
        for (auto& d : deformers_)
            for (size_t i = 0; i < d.selectedMeshes.size(); i++) {
                std::vector<int32_t> g(d.selectedVertices[i].begin(), d.selectedVertices[i].end());
                if (g.empty()) continue;
                if (d.type == 0) ck(gaia_b200_add_rotator(pt.h, 0, g.data(), (int32_t)g.size(), d.center, d.axis, d.angularVelocity, d.endTime), "gaia_b200_add_rotator");
                else ck(gaia_b200_add_translator(pt.h, 0, g.data(), (int32_t)g.size(), d.speed, d.endTime), "gaia_b200_add_translator");
            }
        ck(gaia_b200_finalize(pt.h), "gaia_b200_finalize");
    });
    std::vector<std::array<unsigned char, 16>> ptrs(nGpus_);
    for (int r = 0; r < nGpus_; r++) if (gaia_b200_halo_local_pointers(parts_[r].h, ptrs[r].data()) < 0) throw std::runtime_error(gaia_b200_last_error(parts_[r].h));
    for (int r = 0; r < nGpus_; r++) {
        Part& pt = parts_[r];
        std::vector<int32_t> mine;
        for (int32_t v = 0; v < nV; v++) if (owner[v] == r) mine.push_back(v);
        std::vector<int32_t> mv(2 * mine.size(), 0), remote(std::max<size_t>(mine.size(), 1));
        for (size_t k = 0; k < mine.size(); k++) mv[2 * k + 1] = mine[k];
        for (int q = 0; q < nGpus_; q++) {
            if (q == r) continue;
            if (gaia_b200_get_sorted_ids(parts_[q].h, 0, mine.data(), (int32_t)mine.size(), remote.data()) < 0) throw std::runtime_error(gaia_b200_last_error(parts_[q].h));
            if (gaia_b200_halo_add_peer_pointers(pt.h, q, parts_[q].device, ptrs[q].data(), (int32_t)mine.size(), mv.data(), remote.data()) < 0)
                throw std::runtime_error(gaia_b200_last_error(pt.h));
        }
    }
    for (auto& pt : parts_) if (gaia_b200_halo_finalize(pt.h) < 0) throw std::runtime_error(gaia_b200_last_error(pt.h));
}

void VBDPhysicsB200MultiGPU::runStep() {
    // every partition enqueues (and, where its step needs the host, waits for) its own frame: the colour launches of the
    // partitions wait on each other's epochs, so they must all be in flight together
    forAllParts([&](int r) {
        if (gaia_b200_run_frames(parts_[r].h, 1) < 0) throw std::runtime_error(gaia_b200_last_error(parts_[r].h));
        if (gaia_b200_synchronize(parts_[r].h) < 0) throw std::runtime_error(gaia_b200_last_error(parts_[r].h));
    });
    frameId_ = gaia_b200_frame_id(parts_[0].h);
}

void VBDPhysicsB200MultiGPU::fetchSurfaceTopology() {
    const HostMeshData& hm = meshCache_[models_[0].meshCacheId];
    const int32_t nV = (int32_t)(hm.verts.size() / 3), nT = (int32_t)(hm.tets.size() / 4);
    int32_t nSV = 0, nSF = 0;
    check(gaia_b200_tet_surface(nV, nT, hm.tets.data(), &nSV, nullptr, &nSF, nullptr), "gaia_b200_tet_surface");
    surfIds_.resize(nSV);
    std::vector<int32_t> faces(3 * (size_t)nSF), toSurf(nV, -1);
    check(gaia_b200_tet_surface(nV, nT, hm.tets.data(), &nSV, surfIds_.data(), &nSF, faces.data()), "gaia_b200_tet_surface");
    for (int i = 0; i < nSV; i++) toSurf[surfIds_[i]] = i;
    surfFaces_.assign(1, {}); surfOffset_.assign(2, 0);
    surfFaces_[0].resize(faces.size());
    for (size_t i = 0; i < faces.size(); i++) surfFaces_[0][i] = toSurf[faces[i]];
    surfOffset_[1] = nSV;
}

void VBDPhysicsB200MultiGPU::writeOutputs(const std::string& outFolder, int frameId) {
    if (surfFaces_.empty()) fetchSurfaceTopology();
    forAllParts([&](int r) {      // every partition contributes the vertices it owns
        Part& pt = parts_[r];
        std::vector<float> loc(3 * pt.localVerts.size());
        if (gaia_b200_get_positions(pt.h, 0, loc.data()) < 0) throw std::runtime_error(gaia_b200_last_error(pt.h));
        for (size_t i = 0; i < pt.localVerts.size(); i++)
            if (pt.owned[i]) std::memcpy(&global_[3 * (size_t)pt.localVerts[i]], &loc[3 * i], 12);
    });
    surface_.resize(3 * surfIds_.size());
    for (size_t i = 0; i < surfIds_.size(); i++) std::memcpy(&surface_[3 * i], &global_[3 * (size_t)surfIds_[i]], 12);
    writeFrameFiles(outFolder, frameId, surface_.data());
}

void VBDPhysicsB200MultiGPU::simulate() {
    if (run_.saveOutputs) {
        createFolder(outputFolder_);
        writeOutputs(outputFolder_, frameId_);
    }
    while (frameId_ < run_.numFrames) {
        const auto t0 = std::chrono::high_resolution_clock::now();
        runStep();
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - t0).count();
        if (run_.saveOutputs) writeOutputs(outputFolder_, frameId_);
        if (run_.debugVerboseLvl >= 1) std::cout << "Frame " << frameId_ << " completed on " << nGpus_ << " GPUs, Time consumption: " << ms << "\n";
    }
}

}  // namespace gaia_b200
