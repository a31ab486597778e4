// Host adapter over the C ABI (see VBDPhysicsB200.h for the reference entry points each method mirrors).
#include "VBDPhysicsB200.h"

#include <sys/stat.h>

#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <stdexcept>

#include "host_util.h"
#include "json_min.h"

namespace gaia_b200 {

using jsonmin::Value;

VBDPhysicsB200::~VBDPhysicsB200() {
    try { finishWriter(); } catch (...) {}
    gaia_b200_destroy(h_);
}

void VBDPhysicsB200::check(int rc, const char* what) {
    if (rc < 0) throw std::runtime_error(std::string(what) + " failed (" + std::to_string(rc) + "): " + gaia_b200_last_error(h_));
}

std::string VBDPhysicsB200::resolve(const std::string& p) const {
    std::string r = p;
    if (!repoRoot_.empty()) r = replaceAll(r, "${REPO_ROOT}", repoRoot_);
    return replaceAll(r, "\\", "/");       // checked-in experiment files carry Windows separators
}

// MeshFrame2 `.t`: "Vertex <id> x y z" / "Tet <id> v0 v1 v2 v3"; ids are remapped to 0-based in read order
// (3P/MeshFrame2/MeshFrame/TetMesh/TMeshStatic.h:95-130)
void VBDPhysicsB200::loadTetMesh(const std::string& path, HostMeshData& out) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open tet mesh " + path);
    out.path = path;
    std::map<long, int32_t> vid;
    std::string tok;
    std::string line;
    while (std::getline(f, line)) {
        if (line.empty()) continue;
        const char* c = line.c_str();
        if (!std::strncmp(c, "Vertex", 6)) {
            long id; float x, y, z;
            if (std::sscanf(c + 6, "%ld %f %f %f", &id, &x, &y, &z) != 4) throw std::runtime_error("bad Vertex line in " + path);
            vid[id] = (int32_t)(out.verts.size() / 3);
            out.verts.push_back(x); out.verts.push_back(y); out.verts.push_back(z);
        } else if (!std::strncmp(c, "Tet", 3)) {
            long id, v[4];
            if (std::sscanf(c + 3, "%ld %ld %ld %ld %ld", &id, &v[0], &v[1], &v[2], &v[3]) != 5) throw std::runtime_error("bad Tet line in " + path);
            for (int k = 0; k < 4; k++) {
                auto it = vid.find(v[k]);
                if (it == vid.end()) throw std::runtime_error("Tet references an unknown vertex in " + path);
                out.tets.push_back(it->second);
            }
        }
    }
    if (out.verts.empty() || out.tets.empty()) throw std::runtime_error("empty tet mesh " + path);
}

std::vector<std::vector<int32_t>> loadVertexColoring(const std::string& path) {
    const Value j = jsonmin::parse_file(path);
    if (!j.is_array()) throw std::runtime_error("colouring file is not a list of lists: " + path);
    std::vector<std::vector<int32_t>> out;
    for (auto& c : j.arr) {
        out.emplace_back();
        for (auto& v : c.arr) out.back().push_back((int32_t)v.as_double());
    }
    return out;
}

void VBDPhysicsB200::loadRunningparameters(std::string inModelInputFile, std::string inParameterFile, std::string outFolder) {
    inModelInputFile = resolve(inModelInputFile); inParameterFile = resolve(inParameterFile); outputFolder_ = resolve(outFolder);
    gaia_b200_default_physics_params(&phys_);
    gaia_b200_default_collision_params(&col_);
    const Value params = jsonmin::parse_file(inParameterFile);
    const Value modelsFile = jsonmin::parse_file(inModelInputFile);
    if (const Value* p = params.find("PhysicsParams")) {
        p->get_vec3("gravity", phys_.gravity);
        if (const Value* wb = p->find("worldBounds")) {
            if (wb->is_array() && wb->size() == 2)
                for (int d = 0; d < 3; d++) { phys_.worldBoundsMin[d] = (float)(*wb)[0][d].as_double(); phys_.worldBoundsMax[d] = (float)(*wb)[1][d].as_double(); }
        }
        p->get("timeStep", phys_.timeStep); p->get("numSubsteps", phys_.numSubsteps); p->get("iterations", phys_.iterations);
        bool b;
        if (p->get("usePlaneGround", b)) phys_.usePlaneGround = b;
        if (p->get("useBowlGround", b)) phys_.useBowlGround = b;
        p->get_vec3("bowlCenter", bowlCenter_); p->get("bowlRadius", bowlRadius_); p->get("bowlCap", bowlCap_);
        if (p->get("checkAndUpdateWorldBounds", b)) phys_.checkAndUpdateWorldBounds = b;
        p->get("boundaryFrictionDynamic", phys_.boundaryFrictionDynamic);
        p->get("dcdTetMeshSceneBVHRebuildSteps", phys_.dcdTetMeshSceneBVHRebuildSteps);
        p->get("dcdSurfaceSceneBVHRebuildSteps", phys_.dcdSurfaceSceneBVHRebuildSteps);
        p->get("ccdBVHRebuildSteps", phys_.ccdBVHRebuildSteps);
        p->get("stepSize", phys_.stepSize); p->get("stepSizeGD", phys_.stepSizeGD);
        if (p->get("associateGravityWithInertia", b)) phys_.associateGravityWithInertia = b;
        p->get("collisionStiffness", phys_.collisionStiffness);
        p->get("intermediateCollisionIterations", phys_.intermediateCollisionIterations);
        p->get("boundaryCollisionStiffness", phys_.boundaryCollisionStiffness);
        p->get("boundaryFrictionEpsV", phys_.boundaryFrictionEpsV);
        p->get("frictionStartIter", phys_.frictionStartIter);
        if (p->get("useAccelerator", b)) phys_.useAccelerator = b;
        p->get("acceleratorRho", phys_.acceleratorRho);
        p->get("collisionOffHeight", phys_.collisionOffHeight);
        // host loop
        p->get("numFrames", run_.numFrames);
        p->get("saveOutputs", run_.saveOutputs); p->get("saveAllModelsTogether", run_.saveAllModelsTogether);
        p->get("outputExt", run_.outputExt); p->get("binaryModeVisualizationSteps", run_.binaryModeVisualizationSteps);
        p->get("outputRecoveryState", run_.outputRecoveryState); p->get("outputRecoveryStateStep", run_.outputRecoveryStateStep);
        p->get("outputRecoveryStateBinary", run_.outputRecoveryStateBinary); p->get("outputStatistics", run_.outputStatistics);
        p->get("debugVerboseLvl", run_.debugVerboseLvl);
        p->get("applyLocalLineSearch", run_.applyLocalLineSearch); p->get("tau", run_.tau); p->get("tau_GD", run_.tauGD);
        p->get("backtracingLineSearchMaxIters", run_.backtracingLineSearchMaxIters);
        int solType = 1; bool useNewton = false, useGD = false;
        p->get("collisionSolutionType", solType); p->get("useNewton", useNewton); p->get("useGDSolver", useGD);
        if (useNewton || useGD) throw std::runtime_error("useNewton / useGDSolver select comparison solvers outside the VBD hot path");
        if (solType != 1) std::cerr << "[gaia_b200] collisionSolutionType " << solType << " requested; the hybrid handling (1) is what this path implements\n";
    }
    if (const Value* c = params.find("CollisionParams")) {
        bool b;
#define COLFLAG(name) if (c->get(#name, b)) col_.name = b
        COLFLAG(allowCCD); COLFLAG(allowDCD); COLFLAG(checkFeasibleRegion); COLFLAG(checkTetTraverse); COLFLAG(handleSelfCollision);
        COLFLAG(stopTraversingAfterPassingQueryPoint); COLFLAG(tetrahedralTraverseForNonSelfIntersection); COLFLAG(useStaticTraverse);
        COLFLAG(loopLessTraverse); COLFLAG(shiftQueryPointToCenter);
#undef COLFLAG
        c->get("centerShiftLevel", col_.centerShiftLevel); c->get("maxNumberOfBVHQuery", col_.maxNumberOfBVHQuery);
        c->get("feasibleRegionEpsilon", col_.feasibleRegionEpsilon); c->get("rayTriIntersectionEPSILON", col_.rayTriIntersectionEPSILON);
        c->get("maxSearchDistanceMultiplier", col_.maxSearchDistanceMultiplier);
    }
    if (const Value* ds = params.find("Deformers")) {
        for (auto& d : ds->arr) {
            DeformerInput in;
            std::string name;
            d.get("DeformerName", name);
            if (name == "Rotator") { in.type = 0; d.get_vec3("center", in.center); d.get_vec3("axis", in.axis); d.get("angularVelocity", in.angularVelocity); d.get("rotationEndTime", in.endTime); }
            else if (name == "Translator") { in.type = 1; d.get_vec3("speed", in.speed); d.get("deformationEndTime", in.endTime); }
            else throw std::runtime_error("Unrecognized deformer: " + name);
            d.get_ints("selectedMeshes", in.selectedMeshes);
            if (const Value* sv = d.find("selectedVertices"))
                for (auto& l : sv->arr) { in.selectedVertices.emplace_back(); for (auto& v : l.arr) in.selectedVertices.back().push_back((int)v.as_double()); }
            if (in.selectedVertices.size() != in.selectedMeshes.size()) throw std::runtime_error("deformer: selectedMeshes / selectedVertices size mismatch");
            deformers_.push_back(in);
        }
    }
    const Value* ml = modelsFile.find("Models");
    if (!ml || !ml->is_array()) throw std::runtime_error("Models.json has no \"Models\" list");
    for (auto& m : ml->arr) {
        ModelInput in;
        gaia_b200_default_object_params(&in.obj);
        m.get("materialName", in.materialName);
        if (in.materialName != "NeoHookean") throw std::runtime_error("material " + in.materialName + " is not supported by the VBD path (the reference asserts too, M/VBD/VBDPhysics.cpp:775-786)");
        m.get("path", in.path); m.get("verticesColoringCategoriesPath", in.verticesColoringCategoriesPath);
        in.path = resolve(in.path); in.verticesColoringCategoriesPath = resolve(in.verticesColoringCategoriesPath);
        m.get_vec3("scale", in.obj.scale); m.get_vec3("rotation", in.obj.rotation); m.get_vec3("translation", in.obj.translation);
        m.get_vec3("initialVelocity", in.obj.initialVelocity);
        float tbs[3] = {0, 0, 0};
        if (m.get_vec3("translationBeforeScaling", tbs) && (tbs[0] != 0 || tbs[1] != 0 || tbs[2] != 0))
            std::cerr << "[gaia_b200] translationBeforeScaling is parsed but never applied by the reference either\n";
        m.get("density", in.obj.density);
        bool b;
        if (m.get("hasNoGravZone", b)) in.obj.hasNoGravZone = b;
        m.get("noGravZoneThreshold", in.obj.noGravZoneThreshold); m.get("maxVelocityMagnitude", in.obj.maxVelocityMagnitude);
        m.get("frameToAppear", in.obj.frameToAppear);
        m.get("frictionDynamic", in.obj.frictionDynamic); m.get("frictionEpsV", in.obj.frictionEpsV);
        m.get("exponentialVelDamping", in.obj.exponentialVelDamping); m.get("constantVelDamping", in.obj.constantVelDamping);
        m.get("initializationType", in.obj.initializationType);
        m.get("miu", in.obj.miu); m.get("lmbd", in.obj.lmbd);
        m.get("dampingHydrostatic", in.obj.dampingHydrostatic); m.get("dampingDeviatoric", in.obj.dampingDeviatoric);
        std::vector<int> fp;
        m.get_ints("fixedPoints", fp);
        in.fixedPoints.assign(fp.begin(), fp.end());
        models_.push_back(in);
    }
    loaded_ = true;
}

void VBDPhysicsB200::loadMeshes(int64_t* nVertsTotal, int64_t* nTetsTotal) {
    if (!loaded_) throw std::runtime_error("loadMeshes() before loadRunningparameters()");
    std::map<std::string, int> byPath;          // meshes are loaded once per path (BasePhysicsFramework.cpp:139-160)
    int64_t nv = 0, nt = 0;
    for (auto& m : models_) {
        auto it = byPath.find(m.path);
        if (it == byPath.end()) {
            meshCache_.emplace_back();
            loadTetMesh(m.path, meshCache_.back());
            it = byPath.emplace(m.path, (int)meshCache_.size() - 1).first;
        }
        m.meshCacheId = it->second;
        if (!colourings_.count(m.verticesColoringCategoriesPath))
            colourings_.emplace(m.verticesColoringCategoriesPath, loadVertexColoring(m.verticesColoringCategoriesPath));
        nv += (int64_t)meshCache_[m.meshCacheId].verts.size() / 3; nt += (int64_t)meshCache_[m.meshCacheId].tets.size() / 4;
    }
    if (nVertsTotal) *nVertsTotal = nv;
    if (nTetsTotal) *nTetsTotal = nt;
}

void VBDPhysicsB200::initialize() {
    if (!loaded_) throw std::runtime_error("initialize() before loadRunningparameters()");
    check(gaia_b200_create(&phys_, &col_, device_, &h_), "gaia_b200_create");
    if (phys_.useBowlGround) check(gaia_b200_set_bowl_ground(h_, bowlCenter_, bowlRadius_, bowlCap_ ? 1 : 0), "gaia_b200_set_bowl_ground");
    if (meshCache_.empty()) loadMeshes(nullptr, nullptr);
    for (auto& m : models_) {
        auto ct = colourings_.find(m.verticesColoringCategoriesPath);
        const HostMeshData& hm = meshCache_[m.meshCacheId];
        std::vector<int32_t> offs(1, 0), flat;
        for (auto& c : ct->second) { flat.insert(flat.end(), c.begin(), c.end()); offs.push_back((int32_t)flat.size()); }
        const int rc = gaia_b200_add_tet_mesh(h_, (int32_t)(hm.verts.size() / 3), (int32_t)(hm.tets.size() / 4), hm.verts.data(), hm.tets.data(), &m.obj,
                                              (int32_t)ct->second.size(), offs.data(), flat.data(), (int32_t)m.fixedPoints.size(),
                                              m.fixedPoints.empty() ? nullptr : m.fixedPoints.data());
        check(rc, "gaia_b200_add_tet_mesh");
    }
    for (auto& d : deformers_)
        for (size_t i = 0; i < d.selectedMeshes.size(); i++) {
            std::vector<int32_t> v(d.selectedVertices[i].begin(), d.selectedVertices[i].end());
            if (d.type == 0) check(gaia_b200_add_rotator(h_, d.selectedMeshes[i], v.data(), (int32_t)v.size(), d.center, d.axis, d.angularVelocity, d.endTime), "gaia_b200_add_rotator");
            else check(gaia_b200_add_translator(h_, d.selectedMeshes[i], v.data(), (int32_t)v.size(), d.speed, d.endTime), "gaia_b200_add_translator");
        }
    check(gaia_b200_finalize(h_), "gaia_b200_finalize");
    if (run_.applyLocalLineSearch)
        check(gaia_b200_set_line_search(h_, 1, run_.tau, run_.tauGD, run_.backtracingLineSearchMaxIters), "gaia_b200_set_line_search");
    // single-GPU driver: 8 lanes per vertex when every parallel group is small (see gaia_b200_set_solver_lanes)
    if (models_.size() > 1) {
        int32_t largest = 0;
        for (int g = 0; g < gaia_b200_num_parallel_groups(h_); g++) {
            int32_t n = 0;
            check(gaia_b200_get_parallel_group(h_, g, nullptr, &n), "gaia_b200_get_parallel_group");
            largest = n > largest ? n : largest;
        }
        check(gaia_b200_set_solver_lanes(h_, largest < 65536 ? 8 : 4), "gaia_b200_set_solver_lanes");
    }
    frameId_ = 0;
}

void VBDPhysicsB200::runStep() {
    check(gaia_b200_run_frames(h_, 1), "gaia_b200_run_frames");
    frameId_ = gaia_b200_frame_id(h_);
}

void VBDPhysicsB200::simulate() {
    if (run_.saveOutputs) {
        createFolder(outputFolder_);
        if (run_.outputStatistics) createFolder(outputFolder_ + "/Statistics");
        if (run_.outputRecoveryState) createFolder(outputFolder_ + "/RecoveryStates");
        writeOutputs(outputFolder_, frameId_);
    }
    while (frameId_ < run_.numFrames) {
        const auto t0 = std::chrono::high_resolution_clock::now();
        runStep();
        if (run_.saveOutputs) {
            submitFrame(frameId_);       // surface gather + copy enqueued behind the step; the writer thread takes it from there
            if (run_.outputRecoveryState && !(frameId_ % run_.outputRecoveryStateStep)) {
                createFolder(outputFolder_ + "/RecoveryStates");
                saveRecoveryState(outputFolder_ + "/RecoveryStates/" + numbered("A", frameId_, ".json"), run_.outputRecoveryStateBinary);
            }
        }
        if (run_.debugVerboseLvl >= 1) {
            check(gaia_b200_synchronize(h_), "gaia_b200_synchronize");
            const double ms = std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - t0).count();
            std::cout << "Frame " << frameId_ << " completed, Time consumption: " << ms << "\n";
        }
        if (run_.saveOutputs && run_.outputStatistics) writeStatistics(outputFolder_, frameId_);
    }
    finishWriter();
    check(gaia_b200_synchronize(h_), "gaia_b200_synchronize");
}

// ---- asynchronous frame output
void VBDPhysicsB200::submitFrame(int frameId) {
    const int slot = frameId & 1;
    {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return !slotBusy_[slot] || !writerError_.empty(); });      // the writer is done with this ring slot
        if (!writerError_.empty()) throw std::runtime_error("frame writer: " + writerError_);
        slotBusy_[slot] = true;
    }
    check(gaia_b200_surface_snapshot_begin(h_, slot), "gaia_b200_surface_snapshot_begin");
    {
        std::lock_guard<std::mutex> lk(mu_);
        jobs_.emplace_back(slot, frameId);
        if (!writer_.joinable()) writer_ = std::thread([this] { writerLoop(); });
    }
    cv_.notify_all();
}

void VBDPhysicsB200::writerLoop() {
    for (;;) {
        std::pair<int, int> job;
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] { return stopWriter_ || !jobs_.empty(); });
            if (jobs_.empty()) return;
            job = jobs_.front();
            jobs_.pop_front();
        }
        try {
            const float* surf = nullptr;
            int64_t n = 0;
            const int rc = gaia_b200_surface_snapshot_wait(h_, job.first, &surf, &n);
            if (rc < 0) throw std::runtime_error(std::string("gaia_b200_surface_snapshot_wait: ") + gaia_b200_last_error(h_));
            writeFrameFiles(outputFolder_, job.second, surf);
        } catch (const std::exception& e) {
            std::lock_guard<std::mutex> lk(mu_);
            writerError_ = e.what();
        }
        {
            std::lock_guard<std::mutex> lk(mu_);
            slotBusy_[job.first] = false;
        }
        cv_.notify_all();
    }
}

void VBDPhysicsB200::finishWriter() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        stopWriter_ = true;
    }
    cv_.notify_all();
    if (writer_.joinable()) writer_.join();
    stopWriter_ = false;
    if (!writerError_.empty()) { const std::string e = writerError_; writerError_.clear(); throw std::runtime_error("frame writer: " + e); }
}

void saveAsPly(const std::string& filePath, const std::vector<float>& verts, const std::vector<int32_t>& faces) {
    FILE* fp = std::fopen(filePath.c_str(), "w");
    if (!fp) throw std::runtime_error("cannot write " + filePath);
    std::fprintf(fp, "ply\nformat ascii 1.0\ncomment Mocap generated\nelement vertex %d\n", (int)(verts.size() / 3));
    std::fprintf(fp, "property float x\nproperty float y\nproperty float z\n");
    if (!faces.empty()) std::fprintf(fp, "element face %d \nproperty list uchar int vertex_indices\n", (int)(faces.size() / 3));
    std::fprintf(fp, "end_header\n");
    for (size_t i = 0; i < verts.size() / 3; i++) std::fprintf(fp, "%f %f %f\n", verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
    for (size_t i = 0; i < faces.size() / 3; i++) std::fprintf(fp, "3 %d %d %d\n", faces[3 * i], faces[3 * i + 1], faces[3 * i + 2]);
    std::fclose(fp);
}

void VBDPhysicsB200::writeOutputs(const std::string& outFolder, int frameId) {
    // synchronous form (frame 0, callers that write by hand): same snapshot path, waited for here
    finishWriter();
    check(gaia_b200_surface_snapshot_begin(h_, 0), "gaia_b200_surface_snapshot_begin");
    const float* surf = nullptr;
    int64_t n = 0;
    check(gaia_b200_surface_snapshot_wait(h_, 0, &surf, &n), "gaia_b200_surface_snapshot_wait");
    writeFrameFiles(outFolder, frameId, surf);
    if (run_.outputRecoveryState && !(frameId % run_.outputRecoveryStateStep)) {
        createFolder(outFolder + "/RecoveryStates");
        saveRecoveryState(outFolder + "/RecoveryStates/" + numbered("A", frameId, ".json"), run_.outputRecoveryStateBinary);
    }
}

void VBDPhysicsB200::fetchSurfaceTopology() {      // faces as indices into the surface vertex list, per mesh
    const int nM = gaia_b200_num_meshes(h_);
    surfFaces_.assign(nM, {}); surfOffset_.assign(nM + 1, 0);
    for (int m = 0; m < nM; m++) {
        int32_t nV, nT, nSV, nSF;
        check(gaia_b200_mesh_sizes(h_, m, &nV, &nT, &nSV, &nSF), "gaia_b200_mesh_sizes");
        std::vector<int32_t> ids(nSV), faces(3 * (size_t)nSF), toSurf(nV, -1);
        check(gaia_b200_get_surface_vertex_ids(h_, m, ids.data()), "gaia_b200_get_surface_vertex_ids");
        check(gaia_b200_get_surface_faces(h_, m, faces.data()), "gaia_b200_get_surface_faces");
        for (int i = 0; i < nSV; i++) toSurf[ids[i]] = i;
        surfFaces_[m].resize(faces.size());
        for (size_t i = 0; i < faces.size(); i++) surfFaces_[m][i] = toSurf[faces[i]];
        surfOffset_[m + 1] = surfOffset_[m] + nSV;
    }
}

// writeAllToBinary / writeAllToPLY (M/IO/FileIO.cpp:52-185) from one snapshot: surface vertices of all meshes, concatenated
void VBDPhysicsB200::writeFrameFiles(const std::string& outFolder, int frameId, const float* surfAll) {
    const int nM = (int)models_.size();
    if ((int)surfFaces_.size() != nM) fetchSurfaceTopology();      // static topology, fetched once
    auto verts = [&](int m) { return std::vector<float>(surfAll + 3 * surfOffset_[m], surfAll + 3 * surfOffset_[m + 1]); };
    auto writePly = [&](const std::string& file, bool together, bool addModelName) {
        if (together) {
            std::vector<float> v(surfAll, surfAll + 3 * surfOffset_[nM]);
            std::vector<int32_t> f;
            for (int m = 0; m < nM; m++) for (int32_t id : surfFaces_[m]) f.push_back(id + (int32_t)surfOffset_[m]);
            saveAsPly(file, v, f);
        } else {
            const size_t dot = file.find_last_of('.');
            const std::string stem = file.substr(0, dot), ext = file.substr(dot);
            for (int m = 0; m < nM; m++) {
                std::string name = stem + numbered("_Model_", m, "", 6);
                if (addModelName) name += "_" + baseName(models_[m].path);
                saveAsPly(name + ext, verts(m), surfFaces_[m]);
            }
        }
    };
    const std::string outFile = outFolder + "/" + numbered("A", frameId, ".") + run_.outputExt;
    if (run_.outputExt == "bin") {
        if (frameId == 0) {
            createFolder(outFolder + "/TemplateMesh");
            writePly(outFolder + "/TemplateMesh/TemplateMesh.ply", false, true);
        }
        std::ofstream out(outFile, std::ios::out | std::ios::binary);
        out.write(reinterpret_cast<const char*>(surfAll), (std::streamsize)(3 * surfOffset_[nM] * sizeof(float)));
        out.close();
        if (!(frameId % run_.binaryModeVisualizationSteps)) writePly(outFolder + "/" + numbered("A", frameId, ".ply"), run_.saveAllModelsTogether, run_.saveAllModelsTogether);
    } else if (run_.outputExt == "ply") {
        writePly(outFile, run_.saveAllModelsTogether, run_.saveAllModelsTogether);
    } else {
        std::cout << "[Error] Unsupported output EXT name: " << run_.outputExt << std::endl;
    }
}

void VBDPhysicsB200::saveRecoveryState(const std::string& file, bool binary) {
    const int nM = gaia_b200_num_meshes(h_);
    std::ofstream j(file);
    if (!j) throw std::runtime_error("cannot write " + file);
    j.precision(17);
    j << "{\n  \"frameId\": " << frameId_ << ",\n  \"curTime\": " << gaia_b200_cur_time(h_) << ",\n";
    std::vector<std::vector<float>> pos(nM), vel(nM);
    for (int m = 0; m < nM; m++) {
        int32_t nV, nT, nSV, nSF;
        gaia_b200_mesh_sizes(h_, m, &nV, &nT, &nSV, &nSF);
        pos[m].resize(3 * (size_t)nV); vel[m].resize(3 * (size_t)nV);
        check(gaia_b200_get_positions(h_, m, pos[m].data()), "gaia_b200_get_positions");
        check(gaia_b200_get_velocities(h_, m, vel[m].data()), "gaia_b200_get_velocities");
    }
    if (binary) {
        const std::string base = baseName(file), dir = dirName(file);
        std::ofstream velFile(dir + "/" + base + ".vel", std::ios::binary), posFile(dir + "/" + base + ".pos", std::ios::binary);
        j << "  \"format\": \"binary\",\n  \"nMeshes\": " << nM << ",\n  \"meshesState\": {\n    \"velocities_file\": \"" << base << ".vel\",\n"
          << "    \"position_file\": \"" << base << ".pos\",\n    \"nVerts\": [";
        for (int m = 0; m < nM; m++) {
            velFile.write(reinterpret_cast<const char*>(vel[m].data()), (std::streamsize)(vel[m].size() * 4));
            posFile.write(reinterpret_cast<const char*>(pos[m].data()), (std::streamsize)(pos[m].size() * 4));
            j << (m ? ", " : "") << pos[m].size() / 3;
        }
        j << "]\n  }\n}\n";
    } else {
        j.precision(9);
        j << "  \"format\": \"text\",\n  \"meshesState\": [\n";
        auto dump = [&](const char* key, const std::vector<float>& a) {
            j << "      \"" << key << "\": [";
            for (size_t i = 0; i < a.size() / 3; i++) j << (i ? ", " : "") << "[" << a[3 * i] << ", " << a[3 * i + 1] << ", " << a[3 * i + 2] << "]";
            j << "]";
        };
        for (int m = 0; m < nM; m++) {
            j << "    {\n"; dump("velocities", vel[m]); j << ",\n"; dump("position", pos[m]); j << "\n    }" << (m + 1 < nM ? "," : "") << "\n";
        }
        j << "  ]\n}\n";
    }
}

void VBDPhysicsB200::recoverFromState(const std::string& stateFileIn) {
    const std::string stateFile = resolve(stateFileIn);
    std::cout << "----------------------------------------------------\nRecovering state from:" << stateFile
              << "\n----------------------------------------------------\n";
    const Value j = jsonmin::parse_file(stateFile);
    int frameId = -1; double curTime = 0; std::string format = "text";
    j.get("frameId", frameId); j.get("curTime", curTime); j.get("format", format);
    const int nM = gaia_b200_num_meshes(h_);
    const Value* ms = j.find("meshesState");
    if (!ms) throw std::runtime_error("recovery state has no meshesState");
    if (format == "binary") {
        int nMeshes = 0;
        j.get("nMeshes", nMeshes);
        std::vector<int> nVerts;
        ms->get_ints("nVerts", nVerts);
        std::string vf, pf;
        ms->get("velocities_file", vf); ms->get("position_file", pf);
        std::ifstream velFile(dirName(stateFile) + "/" + vf, std::ios::binary), posFile(dirName(stateFile) + "/" + pf, std::ios::binary);
        if (!velFile || !posFile) throw std::runtime_error("cannot open the .pos/.vel files of " + stateFile);
        for (int m = 0; m < nMeshes && m < nM; m++) {
            int32_t nV, nT, nSV, nSF;
            gaia_b200_mesh_sizes(h_, m, &nV, &nT, &nSV, &nSF);
            if ((int)nVerts.size() <= m || nVerts[m] != nV) throw std::runtime_error("recovery state vertex count mismatch");
            std::vector<float> p(3 * (size_t)nV), v(3 * (size_t)nV);
            velFile.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(v.size() * 4));
            posFile.read(reinterpret_cast<char*>(p.data()), (std::streamsize)(p.size() * 4));
            if (velFile.gcount() != (std::streamsize)(v.size() * 4) || posFile.gcount() != (std::streamsize)(p.size() * 4))
                throw std::runtime_error("recovery state: the .pos / .vel file of " + stateFile + " is shorter than its meshes");
            check(gaia_b200_set_state(h_, m, p.data(), v.data()), "gaia_b200_set_state");
        }
    } else {
        for (size_t m = 0; m < ms->size() && (int)m < nM; m++) {
            const Value& st = (*ms)[m];
            auto flat = [&](const char* key) {
                std::vector<float> out;
                if (const Value* a = st.find(key)) for (auto& p : a->arr) for (int d = 0; d < 3; d++) out.push_back((float)p[d].as_double());
                return out;
            };
            std::vector<float> p = flat("position"), v = flat("velocities");
            int32_t nV, nT, nSV, nSF;
            gaia_b200_mesh_sizes(h_, (int)m, &nV, &nT, &nSV, &nSF);
            if (p.size() != 3 * (size_t)nV || v.size() != 3 * (size_t)nV)      // set_state reads 3 * nV floats from each
                throw std::runtime_error("recovery state vertex count mismatch (mesh " + std::to_string(m) + ": " + std::to_string(p.size() / 3) +
                                         " positions, " + std::to_string(v.size() / 3) + " velocities, " + std::to_string(nV) + " vertices)");
            check(gaia_b200_set_state(h_, (int)m, p.data(), v.data()), "gaia_b200_set_state");
        }
    }
    check(gaia_b200_set_time(h_, frameId, curTime), "gaia_b200_set_time");
    frameId_ = frameId;
}

// per-frame statistics with the reference's timer names (M/Timer/RunningTimeStatistics.h:115-144)
void VBDPhysicsB200::writeStatistics(const std::string& outFolder, int frameId) {
    gaia_stats st;
    check(gaia_b200_get_stats(h_, &st), "gaia_b200_get_stats");
    createFolder(outFolder + "/Statistics");
    std::ofstream j(outFolder + "/Statistics/" + numbered("A", frameId, ".json"));
    j << "{\n  \"timeCsmpInitialStep\": " << st.timeCsmpInitialStep << ",\n  \"timeCsmpMaterialSolve\": " << st.timeCsmpMaterialSolve
      << ",\n  \"timeCsmpUpdateVelocity\": " << st.timeCsmpUpdateVelocity << ",\n  \"timeCsmpUpdatingBVHDCD\": " << st.timeCsmpUpdatingBVHDCD
      << ",\n  \"timeCsmpColDetectDCD\": " << st.timeCsmpColDetectDCD << ",\n  \"timeCsmpUpdatingBVHCCD\": " << st.timeCsmpUpdatingBVHCCD
      << ",\n  \"timeCsmpColDetectCCD\": " << st.timeCsmpColDetectCCD << ",\n  \"timeCsmpFrame\": " << st.timeCsmpFrame << ",\n  \"solveKernelMs\": " << st.solveKernelMs
      << ",\n  \"vertexSolves\": " << st.vertexSolves << ",\n  \"colourSweeps\": " << st.colourSweeps << ",\n  \"substeps\": " << st.substeps
      << ",\n  \"kernelLaunches\": " << st.kernelLaunches << ",\n  \"algorithmicBytes\": " << st.algorithmicBytes
      << ",\n  \"numContactsDCD\": " << st.numContactsDCD << ",\n  \"numContactsCCD\": " << st.numContactsCCD << "\n}\n";
    gaia_b200_reset_stats(h_);
}

}  // namespace gaia_b200
