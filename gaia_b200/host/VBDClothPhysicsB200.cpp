// Cloth host adapter over the cloth C ABI (see VBDClothPhysicsB200.h for the reference entry points each method mirrors).
#include "VBDClothPhysicsB200.h"

#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <stdexcept>

#include "VBDPhysicsB200.h"   // loadVertexColoring, saveAsPly
#include "host_util.h"
#include "json_min.h"

namespace gaia_b200 {

using jsonmin::Value;

VBDClothPhysicsB200::~VBDClothPhysicsB200() { gaia_b200_cloth_destroy(h_); }

void VBDClothPhysicsB200::check(int rc, const char* what) {
    if (rc < 0) throw std::runtime_error(std::string(what) + " failed (" + std::to_string(rc) + "): " + gaia_b200_cloth_last_error(h_));
}

std::string VBDClothPhysicsB200::resolve(const std::string& p) const {
    std::string r = p;
    if (!repoRoot_.empty()) r = replaceAll(r, "${REPO_ROOT}", repoRoot_);
    return replaceAll(r, "\\", "/");
}

// Triangle .obj: "v x y z" and "f a b c" with optional /vt/vn suffixes, 1-based or negative (relative) indices
void loadObj(const std::string& path, HostTriMeshData& out) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open obj mesh " + path);
    out.path = path;
    std::string line;
    while (std::getline(f, line)) {
        const char* c = line.c_str();
        if (c[0] == 'v' && (c[1] == ' ' || c[1] == '\t')) {
            float x, y, z;
            if (std::sscanf(c + 2, "%f %f %f", &x, &y, &z) != 3) throw std::runtime_error("bad v line in " + path);
            out.verts.push_back(x); out.verts.push_back(y); out.verts.push_back(z);
        } else if (c[0] == 'f' && (c[1] == ' ' || c[1] == '\t')) {
            const char* p = c + 2;
            int n = 0;
            long ids[4];
            while (*p) {
                while (*p == ' ' || *p == '\t' || *p == '\r') p++;
                if (!*p) break;
                char* end = nullptr;
                const long id = std::strtol(p, &end, 10);
                if (end == p) throw std::runtime_error("bad f line in " + path);
                if (n < 4) ids[n] = id;
                n++;
                p = end;
                while (*p && *p != ' ' && *p != '\t') p++;     // skip /vt/vn
            }
            if (n != 3) throw std::runtime_error(path + ": only triangle faces are supported");
            const long nV = (long)(out.verts.size() / 3);
            for (int k = 0; k < 3; k++) {
                const long v = ids[k] > 0 ? ids[k] - 1 : nV + ids[k];
                if (v < 0 || v >= nV) throw std::runtime_error("face references an unknown vertex in " + path);
                out.faces.push_back((int32_t)v);
            }
        }
    }
    if (out.verts.empty() || out.faces.empty()) throw std::runtime_error("empty triangle mesh " + path);
}

void VBDClothPhysicsB200::loadRunningparameters(std::string inModelInputFile, std::string inParameterFile, std::string outFolder) {
    inModelInputFile = resolve(inModelInputFile); inParameterFile = resolve(inParameterFile); outputFolder_ = resolve(outFolder);
    gaia_b200_cloth_default_physics_params(&phys_);
    const Value params = jsonmin::parse_file(inParameterFile);
    const Value modelsFile = jsonmin::parse_file(inModelInputFile);
    if (const Value* ds = params.find("Deformers")) {      // loadClothDeformers, M/VBDCloth/VBDClothDeformer.h:202-242
        for (auto& d : ds->arr) {
            ClothDeformerInput in;
            std::string name;
            d.get("DeformerName", name);
            if (name == "Rotator") { in.type = 0; d.get_vec3("center", in.center); d.get_vec3("axis", in.axis); d.get("angularVelocity", in.angularVelocity); d.get("rotationEndTime", in.rotationEndTime); }
            else if (name == "Translator") { in.type = 1; d.get_vec3("speed", in.speed); d.get("deformationStartTime", in.deformationStartTime); d.get("deformationEndTime", in.deformationEndTime); }
            else throw std::runtime_error("Unrecognized deformer: " + name);
            d.get_ints("selectedMeshes", in.selectedMeshes);
            if (const Value* sv = d.find("selectedVertices"))
                for (auto& l : sv->arr) { in.selectedVertices.emplace_back(); for (auto& v : l.arr) in.selectedVertices.back().push_back((int32_t)v.as_double()); }
            if (in.selectedVertices.size() != in.selectedMeshes.size()) throw std::runtime_error("deformer: selectedMeshes / selectedVertices size mismatch");
            deformers_.push_back(in);
        }
    }
    if (const Value* cp = params.find("ColliderParams"))       // createColliderMesh, M/Framework/BaseClothSimPhsicsFramework.cpp:95-114
        if (const Value* cm = cp->find("ColliderMeshes"))
            for (auto& c : cm->arr) {
                ColliderInput in;
                std::string type;
                c.get("colliderType", type);
                if (type != "TriMeshSequence") throw std::runtime_error("Unrecognized collider type: " + type);
                gaia_b200_cloth_default_object_params(&in.obj);
                c.get_vec3("scale", in.obj.scale); c.get_vec3("rotation", in.obj.rotation); c.get_vec3("translation", in.obj.translation);
                c.get("frictionDynamic", in.obj.frictionDynamic); c.get("frictionEpsV", in.obj.frictionEpsV);
                if (const Value* mf = c.find("meshFiles")) for (auto& f : mf->arr) in.meshFiles.push_back(resolve(f.str));
                c.get_ints("keyFrames", in.keyFrames);
                c.get("interpolateSubstep", in.interpolateSubstep); c.get("interpolateIter", in.interpolateIter);
                if (in.keyFrames.size() > in.meshFiles.size()) in.keyFrames.resize(in.meshFiles.size());
                if (!in.keyFrames.empty()) in.meshFiles.resize(in.keyFrames.size());
                if (in.meshFiles.empty()) throw std::runtime_error("collider sequence without mesh files");
                colliders_.push_back(in);
            }
    if (const Value* p = params.find("PhysicsParams")) {
        bool b;
        p->get_vec3("gravity", phys_.gravity);
        if (const Value* wb = p->find("worldBounds"))
            if (wb->is_array() && wb->size() == 2)
                for (int d = 0; d < 3; d++) { phys_.worldBoundsMin[d] = (float)(*wb)[0][d].as_double(); phys_.worldBoundsMax[d] = (float)(*wb)[1][d].as_double(); }
        p->get("timeStep", phys_.timeStep); p->get("numSubsteps", phys_.numSubsteps); p->get("iterations", phys_.iterations);
        p->get("boundaryFrictionDynamic", phys_.boundaryFrictionDynamic);
        p->get("stepSize", phys_.stepSize);
        if (p->get("associateGravityWithInertia", b)) phys_.associateGravityWithInertia = b;
        if (p->get("applyAcceleration", b)) phys_.applyAcceleration = b;
        p->get("accelerationRho", phys_.acceleratorRho);      // the parameter scripts write accelerationRho
        p->get("acceleratorRho", phys_.acceleratorRho);
        if (p->get("handleCollision", b)) phys_.handleCollision = b;
        p->get("contactStiffness", phys_.contactStiffness); p->get("contactRadius", phys_.contactRadius);
        p->get("thickness", phys_.thickness); p->get("conservativeStepRelaxation", phys_.conservativeStepRelaxation);
        p->get("boundaryContactStiffness", phys_.boundaryContactStiffness);
        p->get("boundaryFrictionEpsValue", phys_.boundaryFrictionEpsValue);
        p->get("contactEnergyTypes", phys_.contactEnergyTypes);
        p->get("contactBVHReconstructionSteps", phys_.contactBVHReconstructionSteps);
        p->get("degenerateTriangleThres", phys_.degenerateTriangleThres);
        // host loop
        p->get("numFrames", run_.numFrames);
        p->get("saveOutputs", run_.saveOutputs); p->get("saveAllModelsTogether", run_.saveAllModelsTogether);
        p->get("outputExt", run_.outputExt); p->get("binaryModeVisualizationSteps", run_.binaryModeVisualizationSteps);
        p->get("outputRecoveryState", run_.outputRecoveryState); p->get("outputRecoveryStateStep", run_.outputRecoveryStateStep);
        p->get("outputRecoveryStateBinary", run_.outputRecoveryStateBinary); p->get("outputStatistics", run_.outputStatistics);
        p->get("debugVerboseLvl", run_.debugVerboseLvl);
    }
    if (const Value* cd = params.find("ContactDetectorParams")) cd->get("maxQueryDis", phys_.maxQueryDis);
    // VertexFaceAttachmentConstraintParams::fromJson (M/VBDCloth/VBDTriMeshConstraints.h:13-19), read at VBDClothPhysics.cpp:152-159
    if (const Value* p = params.find("PhysicsParams"))
        if (const Value* a = p->find("VertexFaceAttachmentConstraintParams")) {
            a->get_ints("vertexConstraintVertexInfos", attachments_.vertexInfos);
            a->get_ints("vertexConstraintTargetFacesInfos", attachments_.targetFacesInfos);
            if (const Value* b = a->find("vertexConstraintBarys"))
                if (b->is_array()) for (auto& t : b->arr) { if (!t.is_array() || t.arr.size() != 3) throw std::runtime_error("vertexConstraintBarys holds triples"); for (auto& x : t.arr) attachments_.barys.push_back((float)x.as_double()); }
            if (const Value* k = a->find("vertexConstraintStiffness"))
                if (k->is_array()) for (auto& x : k->arr) attachments_.stiffness.push_back((float)x.as_double());
            const size_t n = attachments_.vertexInfos.size() / 2;
            if (attachments_.vertexInfos.size() % 2 || attachments_.targetFacesInfos.size() != 4 * n || attachments_.barys.size() != 3 * n ||
                (!attachments_.stiffness.empty() && attachments_.stiffness.size() != n))
                throw std::runtime_error("VertexFaceAttachmentConstraintParams: (mesh, vertex) pairs, (mesh, v, v, v) quadruples, one barycentric triple "
                                         "and (optionally) one stiffness per constraint");
        }
    const Value* ml = modelsFile.find("Models");
    if (!ml || !ml->is_array()) throw std::runtime_error("Models.json has no \"Models\" list");
    for (auto& m : ml->arr) {
        ClothModelInput in;
        gaia_b200_cloth_default_object_params(&in.obj);
        m.get("path", in.path); m.get("verticesColoringCategoriesPath", in.verticesColoringCategoriesPath);
        in.path = resolve(in.path); in.verticesColoringCategoriesPath = resolve(in.verticesColoringCategoriesPath);
        m.get_vec3("scale", in.obj.scale); m.get_vec3("rotation", in.obj.rotation); m.get_vec3("translation", in.obj.translation);
        m.get_vec3("initialVelocity", in.initialVelocity);
        bool b;
        m.get("density", in.obj.density);
        if (m.get("hasNoGravZone", b)) in.obj.hasNoGravZone = b;
        m.get("noGravZoneThreshold", in.obj.noGravZoneThreshold); m.get("maxVelocityMagnitude", in.obj.maxVelocityMagnitude);
        m.get("frictionDynamic", in.obj.frictionDynamic); m.get("frictionEpsV", in.obj.frictionEpsV);
        m.get("bendingStiffness", in.obj.bendingStiffness);
        m.get("exponentialVelDamping", in.obj.exponentialVelDamping); m.get("constantVelDamping", in.obj.constantVelDamping);
        m.get("initializationType", in.obj.initializationType);
        m.get("miu", in.obj.miu); m.get("lambda", in.obj.lambda);
        m.get("dampingStVK", in.obj.dampingStVK); m.get("dampingBending", in.obj.dampingBending);
        m.get("sewingPath", in.sewingPath); m.get("filteredPairsPath", in.filteredPairsPath); m.get("sewingStiffness", in.sewingStiffness);
        if (!in.sewingPath.empty()) in.sewingPath = resolve(in.sewingPath);
        if (!in.filteredPairsPath.empty()) in.filteredPairsPath = resolve(in.filteredPairsPath);
        std::vector<int> fp;
        m.get_ints("fixedPoints", fp);
        in.fixedPoints.assign(fp.begin(), fp.end());
        models_.push_back(in);
    }
    loaded_ = true;
}

void VBDClothPhysicsB200::loadMeshes(int64_t* nVertsTotal, int64_t* nFacesTotal) {
    if (!loaded_) throw std::runtime_error("loadMeshes() before loadRunningparameters()");
    std::map<std::string, int> byPath;
    int64_t nv = 0, nf = 0;
    for (auto& m : models_) {
        auto it = byPath.find(m.path);
        if (it == byPath.end()) {
            meshCache_.emplace_back();
            loadObj(m.path, meshCache_.back());
            it = byPath.emplace(m.path, (int)meshCache_.size() - 1).first;
        }
        m.meshCacheId = it->second;
        if (!colourings_.count(m.verticesColoringCategoriesPath))
            colourings_.emplace(m.verticesColoringCategoriesPath, loadVertexColoring(m.verticesColoringCategoriesPath));
        nv += (int64_t)meshCache_[m.meshCacheId].verts.size() / 3; nf += (int64_t)meshCache_[m.meshCacheId].faces.size() / 3;
        if (!m.sewingPath.empty() && m.sewingRatios.empty()) {       // TriMeshFEM::initialize, M/TriMesh/TriMesh.cpp:95-110 (".json" only, as there)
            if (m.sewingPath.size() < 5 || m.sewingPath.substr(m.sewingPath.size() - 5) != ".json") throw std::runtime_error("Unsupported file format: " + m.sewingPath);
            const Value sj = jsonmin::parse_file(m.sewingPath);
            if (const Value* sv = sj.find("vertices"))
                for (auto& t : sv->arr) { if (!t.is_array() || t.arr.size() != 3) throw std::runtime_error("sewing \"vertices\" holds triples"); for (auto& x : t.arr) m.sewingVertices.push_back((int32_t)x.as_double()); }
            if (const Value* sr = sj.find("ratios")) for (auto& x : sr->arr) m.sewingRatios.push_back((float)x.as_double());
            if (m.sewingVertices.size() != 3 * m.sewingRatios.size()) throw std::runtime_error("sewing file: one ratio per vertex triple");
            if (!m.filteredPairsPath.empty()) {       // generate_sewing_filters (M/TriMesh/TriMesh.h:603-621): read only beside a sewing file
                const Value fj = jsonmin::parse_file(m.filteredPairsPath);
                auto pairs = [&](const char* key, std::vector<int32_t>& out) {
                    if (const Value* a = fj.find(key))
                        for (auto& t : a->arr) { if (!t.is_array() || t.arr.size() != 2) throw std::runtime_error(std::string("filtered pairs \"") + key + "\" holds pairs"); for (auto& x : t.arr) out.push_back((int32_t)x.as_double()); }
                };
                pairs("vf", m.filteredVF); pairs("ee", m.filteredEE);
            }
        }
    }
    if (nVertsTotal) *nVertsTotal = nv;
    if (nFacesTotal) *nFacesTotal = nf;
}

void VBDClothPhysicsB200::initialize() {
    if (!loaded_) throw std::runtime_error("initialize() before loadRunningparameters()");
    check(gaia_b200_cloth_create(&phys_, device_, &h_), "gaia_b200_cloth_create");
    if (meshCache_.empty()) loadMeshes(nullptr, nullptr);
    for (auto& m : models_) {
        auto ct = colourings_.find(m.verticesColoringCategoriesPath);
        const HostTriMeshData& hm = meshCache_[m.meshCacheId];
        std::vector<int32_t> offs(1, 0), flat;
        for (auto& c : ct->second) { flat.insert(flat.end(), c.begin(), c.end()); offs.push_back((int32_t)flat.size()); }
        const int rc = gaia_b200_cloth_add_tri_mesh(h_, 1, (int32_t)(hm.verts.size() / 3), (int32_t)(hm.faces.size() / 3), hm.verts.data(), hm.faces.data(),
                                                    &m.obj, (int32_t)ct->second.size(), offs.data(), flat.data(), (int32_t)m.fixedPoints.size(),
                                                    m.fixedPoints.empty() ? nullptr : m.fixedPoints.data());
        check(rc, "gaia_b200_cloth_add_tri_mesh");
        if (!m.sewingRatios.empty()) {
            check(gaia_b200_cloth_set_sewing(h_, rc, (int32_t)m.sewingRatios.size(), m.sewingVertices.data(), m.sewingRatios.data(), m.sewingStiffness), "gaia_b200_cloth_set_sewing");
            if (!m.filteredVF.empty() || !m.filteredEE.empty())
                check(gaia_b200_cloth_set_filtered_pairs(h_, rc, (int32_t)(m.filteredVF.size() / 2), m.filteredVF.data(), (int32_t)(m.filteredEE.size() / 2), m.filteredEE.data()),
                      "gaia_b200_cloth_set_filtered_pairs");
        }
    }
    for (auto& c : colliders_) {       // colliderTriMeshes come after the simulated meshes (triMeshesAll)
        std::vector<float> keys;
        HostTriMeshData first;
        for (size_t k = 0; k < c.meshFiles.size(); k++) {
            HostTriMeshData hm;
            loadObj(c.meshFiles[k], hm);
            if (k == 0) first = hm;
            else if (hm.verts.size() != first.verts.size() || hm.faces.size() != first.faces.size()) throw std::runtime_error("collider key frames must share one topology");
            keys.insert(keys.end(), hm.verts.begin(), hm.verts.end());
        }
        const int id = gaia_b200_cloth_add_tri_mesh(h_, 0, (int32_t)(first.verts.size() / 3), (int32_t)(first.faces.size() / 3), first.verts.data(),
                                                    first.faces.data(), &c.obj, 0, nullptr, nullptr, 0, nullptr);
        check(id, "gaia_b200_cloth_add_tri_mesh (collider)");
        std::vector<int32_t> kf(c.keyFrames.begin(), c.keyFrames.end());
        check(gaia_b200_cloth_set_collider_sequence(h_, id, (int32_t)c.meshFiles.size(), kf.empty() ? nullptr : kf.data(), keys.data(),
                                                    c.interpolateSubstep ? 1 : 0, c.interpolateIter), "gaia_b200_cloth_set_collider_sequence");
    }
    for (auto& d : deformers_)
        for (size_t i = 0; i < d.selectedMeshes.size(); i++) {
            const std::vector<int32_t>& v = d.selectedVertices[i];
            if (d.type == 0) check(gaia_b200_cloth_add_rotator(h_, d.selectedMeshes[i], v.data(), (int32_t)v.size(), d.center, d.axis, d.angularVelocity, d.rotationEndTime), "gaia_b200_cloth_add_rotator");
            else check(gaia_b200_cloth_add_translator(h_, d.selectedMeshes[i], v.data(), (int32_t)v.size(), d.speed, d.deformationStartTime, d.deformationEndTime), "gaia_b200_cloth_add_translator");
        }
    for (size_t i = 0; i < attachments_.vertexInfos.size() / 2; i++) {      // the set-up loop of VBDClothPhysics.cpp:161-220
        const int32_t fv[3] = {attachments_.targetFacesInfos[4 * i + 1], attachments_.targetFacesInfos[4 * i + 2], attachments_.targetFacesInfos[4 * i + 3]};
        check(gaia_b200_cloth_add_attachment(h_, attachments_.vertexInfos[2 * i], attachments_.vertexInfos[2 * i + 1], attachments_.targetFacesInfos[4 * i], fv,
                                             &attachments_.barys[3 * i], attachments_.stiffness.empty() ? 1e4f : attachments_.stiffness[i]),
              "gaia_b200_cloth_add_attachment");
    }
    check(gaia_b200_cloth_finalize(h_), "gaia_b200_cloth_finalize");
    for (size_t i = 0; i < models_.size(); i++) {
        const float* v0 = models_[i].initialVelocity;
        if (v0[0] == 0 && v0[1] == 0 && v0[2] == 0) continue;
        int32_t nV, nF, nE;
        check(gaia_b200_cloth_mesh_sizes(h_, (int)i, &nV, &nF, &nE), "gaia_b200_cloth_mesh_sizes");
        std::vector<float> vel(3 * (size_t)nV);
        for (int32_t k = 0; k < nV; k++) { vel[3 * k] = v0[0]; vel[3 * k + 1] = v0[1]; vel[3 * k + 2] = v0[2]; }
        check(gaia_b200_cloth_set_state(h_, (int)i, nullptr, vel.data()), "gaia_b200_cloth_set_state");
    }
    frameId_ = 0;
    curTime_ = 0.0;
}

void VBDClothPhysicsB200::runStep() {
    check(gaia_b200_cloth_run_frames(h_, 1), "gaia_b200_cloth_run_frames");
    frameId_ = gaia_b200_cloth_frame_id(h_);
    curTime_ = gaia_b200_cloth_cur_time(h_);
}

void VBDClothPhysicsB200::simulate() {
    if (run_.saveOutputs) {
        createFolder(outputFolder_);
        if (run_.outputStatistics) createFolder(outputFolder_ + "/Statistics");
        if (run_.outputRecoveryState) createFolder(outputFolder_ + "/RecoveryStates");
        writeOutputs(outputFolder_, frameId_);
    }
    while (frameId_ < run_.numFrames) {
        const auto t0 = std::chrono::high_resolution_clock::now();
        runStep();
        check(gaia_b200_cloth_synchronize(h_), "gaia_b200_cloth_synchronize");
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - t0).count();
        if (run_.saveOutputs) writeOutputs(outputFolder_, frameId_);
        if (run_.debugVerboseLvl >= 1) std::cout << "Frame " << frameId_ << " completed, Time consumption: " << ms << "\n";
        if (run_.saveOutputs && run_.outputStatistics) writeStatistics(outputFolder_, frameId_);
    }
}

// writeAllToPLY / writeAllToBinary over baseTriMeshesForSimulation: all vertices and the mesh's own triangles
void VBDClothPhysicsB200::writeOutputs(const std::string& outFolder, int frameId) {
    const int nM = (int)models_.size();      // baseTriMeshesForSimulation: colliders are not written
    std::vector<std::vector<float>> pos(nM);
    for (int m = 0; m < nM; m++) {
        int32_t nV, nF, nE;
        check(gaia_b200_cloth_mesh_sizes(h_, m, &nV, &nF, &nE), "gaia_b200_cloth_mesh_sizes");
        pos[m].resize(3 * (size_t)nV);
        check(gaia_b200_cloth_get_positions(h_, m, pos[m].data()), "gaia_b200_cloth_get_positions");
    }
    auto writePly = [&](const std::string& file, bool together, bool addModelName) {
        if (together) {
            std::vector<float> v; std::vector<int32_t> f;
            int base = 0;
            for (int m = 0; m < nM; m++) {
                v.insert(v.end(), pos[m].begin(), pos[m].end());
                for (int32_t id : meshCache_[models_[m].meshCacheId].faces) f.push_back(id + base);
                base += (int)(pos[m].size() / 3);
            }
            saveAsPly(file, v, f);
        } else {
            const size_t dot = file.find_last_of('.');
            const std::string stem = file.substr(0, dot), ext = file.substr(dot);
            for (int m = 0; m < nM; m++) {
                std::string name = stem + numbered("_Model_", m, "", 6);
                if (addModelName) name += "_" + baseName(models_[m].path);
                saveAsPly(name + ext, pos[m], meshCache_[models_[m].meshCacheId].faces);
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
        for (int m = 0; m < nM; m++) out.write(reinterpret_cast<const char*>(pos[m].data()), (std::streamsize)(pos[m].size() * sizeof(float)));
        out.close();
        if (!(frameId % run_.binaryModeVisualizationSteps)) writePly(outFolder + "/" + numbered("A", frameId, ".ply"), run_.saveAllModelsTogether, run_.saveAllModelsTogether);
    } else if (run_.outputExt == "ply") {
        writePly(outFile, run_.saveAllModelsTogether, false);
    } else {
        std::cout << "[Error] Unsupported output EXT name: " << run_.outputExt << std::endl;
        return;
    }
    if (run_.outputRecoveryState && !(frameId % run_.outputRecoveryStateStep)) {
        createFolder(outFolder + "/RecoveryStates");
        saveRecoveryState(outFolder + "/RecoveryStates/" + numbered("A", frameId, ".json"), run_.outputRecoveryStateBinary);
    }
}

void VBDClothPhysicsB200::saveRecoveryState(const std::string& file, bool binary) {
    const int nM = (int)models_.size();      // baseTriMeshesForSimulation: colliders are not written
    std::ofstream j(file);
    if (!j) throw std::runtime_error("cannot write " + file);
    j.precision(17);
    j << "{\n  \"frameId\": " << frameId_ << ",\n  \"curTime\": " << curTime_ << ",\n";
    std::vector<std::vector<float>> pos(nM), vel(nM);
    for (int m = 0; m < nM; m++) {
        int32_t nV, nF, nE;
        gaia_b200_cloth_mesh_sizes(h_, m, &nV, &nF, &nE);
        pos[m].resize(3 * (size_t)nV); vel[m].resize(3 * (size_t)nV);
        check(gaia_b200_cloth_get_positions(h_, m, pos[m].data()), "gaia_b200_cloth_get_positions");
        check(gaia_b200_cloth_get_velocities(h_, m, vel[m].data()), "gaia_b200_cloth_get_velocities");
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

void VBDClothPhysicsB200::recoverFromState(const std::string& stateFileIn) {
    const std::string stateFile = resolve(stateFileIn);
    std::cout << "----------------------------------------------------\nRecovering state from:" << stateFile
              << "\n----------------------------------------------------\n";
    const Value j = jsonmin::parse_file(stateFile);
    int frameId = -1; double curTime = 0; std::string format = "text";
    j.get("frameId", frameId); j.get("curTime", curTime); j.get("format", format);
    const int nM = (int)models_.size();      // baseTriMeshesForSimulation: colliders are not written
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
            int32_t nV, nF, nE;
            gaia_b200_cloth_mesh_sizes(h_, m, &nV, &nF, &nE);
            if (m >= (int)nVerts.size() || nVerts[m] != nV) throw std::runtime_error("recovery state vertex count mismatch");
            std::vector<float> p(3 * (size_t)nV), v(3 * (size_t)nV);
            velFile.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(v.size() * 4));
            posFile.read(reinterpret_cast<char*>(p.data()), (std::streamsize)(p.size() * 4));
            check(gaia_b200_cloth_set_state(h_, m, p.data(), v.data()), "gaia_b200_cloth_set_state");
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
            int32_t nV, nF, nE;
            gaia_b200_cloth_mesh_sizes(h_, (int)m, &nV, &nF, &nE);
            if (p.size() != 3 * (size_t)nV || v.size() != 3 * (size_t)nV) throw std::runtime_error("recovery state vertex count mismatch");
            check(gaia_b200_cloth_set_state(h_, (int)m, p.data(), v.data()), "gaia_b200_cloth_set_state");
        }
    }
    check(gaia_b200_cloth_set_time(h_, frameId, curTime), "gaia_b200_cloth_set_time");
    frameId_ = frameId;
    curTime_ = curTime;
}

// per-frame statistics with the reference's timer names (M/Timer/RunningTimeStatistics.h)
void VBDClothPhysicsB200::writeStatistics(const std::string& outFolder, int frameId) {
    gaia_cloth_stats st;
    check(gaia_b200_cloth_get_stats(h_, &st), "gaia_b200_cloth_get_stats");
    createFolder(outFolder + "/Statistics");
    std::ofstream j(outFolder + "/Statistics/" + numbered("A", frameId, ".json"));
    j << "{\n  \"timeCsmpUpdatingCollisionInfoDCD\": " << st.timeCsmpUpdatingCollisionInfoDCD << ",\n  \"timeCsmpInitialStep\": " << st.timeCsmpInitialStep
      << ",\n  \"timeCsmpMaterialSolve\": " << st.timeCsmpMaterialSolve << ",\n  \"solveKernelMs\": " << st.solveKernelMs
      << ",\n  \"vertexSolves\": " << st.vertexSolves << ",\n  \"colourSweeps\": " << st.colourSweeps << ",\n  \"substeps\": " << st.substeps
      << ",\n  \"kernelLaunches\": " << st.kernelLaunches << ",\n  \"algorithmicBytes\": " << st.algorithmicBytes
      << ",\n  \"numberCollisionDetection\": " << st.numberCollisionDetection << ",\n  \"numContactsVF\": " << st.numContactsVF
      << ",\n  \"numContactsEE\": " << st.numContactsEE << "\n}\n";
    gaia_b200_cloth_reset_stats(h_);
}

}  // namespace gaia_b200
