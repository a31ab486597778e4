// VBDClothPhysicsB200 — C++ host adapter with the reference's cloth framework entry points on top of the cloth C ABI of
// libgaia_b200.so (include/gaia_b200_cloth.h). A Gaia-style cloth main (Simulator/VBDCloth/main.cpp) keeps its shape:
//
//     VBDClothPhysicsB200 physics;
//     physics.loadRunningparameters(modelsJson, parametersJson, outFolder);
//     physics.initialize();
//     if (!recoveryStateFile.empty()) physics.recoverFromState(recoveryStateFile);
//     physics.simulate();
//
// Mirrors (names, argument meaning, file formats):
//   BaseClothPhsicsFramework::parseRunningParameters / initialize       M/Framework/BaseClothSimPhsicsFramework.cpp:17-78
//   BaseClothPhsicsFramework::writeOutputs / recoverFromState           M/Framework/BaseClothSimPhsicsFramework.cpp:128-197
//   ClothPhysicsState (text and binary recovery states)                 M/Framework/BaseClothSimPhsicsFramework.h:65-117
//   VBDClothSimulationFramework::runStep (= runStep_CDEveryIter)        M/VBDCloth/VBDClothPhysics.cpp:389-518
//   TriMeshFEM::loadObj (triangle .obj)                                 M/TriMesh/TriMesh.cpp:15-111
// Models are the simulated meshes (baseTriMeshesForSimulation); a collider is a model whose vertices are all fixed, as
// S03_MultiLayerClothOnCollider.py writes it. `Deformers` (Rotator / Translator, M/VBDCloth/VBDClothDeformer.h) are passed
// to the library, and so are `ColliderParams.ColliderMeshes` of type "TriMeshSequence" (key-frame .obj files of one topology,
// M/SpatialQuery/ColiiderTriMeshSequence.h); other collider types are refused loudly.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "gaia_b200_cloth.h"

namespace gaia_b200 {

struct ClothModelInput {
    std::string path, verticesColoringCategoriesPath;
    gaia_cloth_object_params obj;
    std::vector<int32_t> fixedPoints;
    float initialVelocity[3] = {0, 0, 0};
    int meshCacheId = -1;
    // ObjectParams::sewingPath / ::sewingStiffness (M/TriMesh/TriMesh.cpp:95-110, 458): {"vertices": [[v0, v1, v2], ...], "ratios": [...]}
    std::string sewingPath, filteredPairsPath;
    float sewingStiffness = 0.f;
    std::vector<int32_t> sewingVertices;     // 3 per seam, read by loadMeshes()
    std::vector<float> sewingRatios;
    std::vector<int32_t> filteredVF, filteredEE;   // ObjectParams::filteredPairsPath {"vf": [[v, f], ...], "ee": [[e, e'], ...]}, 2 ints per pair
};

// PhysicsParams.VertexFaceAttachmentConstraintParams (M/VBDCloth/VBDTriMeshConstraints.h:6-29)
struct ClothAttachmentInput {
    std::vector<int> vertexInfos, targetFacesInfos;      // (mesh, vertex) * N, (mesh or -1, three face vertices) * N
    std::vector<float> barys, stiffness;                 // 3 * N; N or empty (the reference's default 1e4)
};

struct ClothDeformerInput {
    int type = 0;   // 0 Rotator, 1 Translator
    float center[3] = {0, 0, 0}, axis[3] = {0, 0, 0}, speed[3] = {0, 0, 0};
    double angularVelocity = 0, rotationEndTime = -1;
    float deformationStartTime = 0.f, deformationEndTime = -1.f;
    std::vector<int> selectedMeshes;
    std::vector<std::vector<int32_t>> selectedVertices;
};

struct ColliderInput {
    gaia_cloth_object_params obj;
    std::vector<std::string> meshFiles;
    std::vector<int> keyFrames;
    bool interpolateSubstep = true;
    int interpolateIter = 1;
};

struct HostTriMeshData {
    std::string path;
    std::vector<float> verts;      // 3*nV
    std::vector<int32_t> faces;    // 3*nF
};

struct ClothHostRunParams {
    int numFrames = 5000;
    bool saveOutputs = true;
    bool saveAllModelsTogether = true;
    std::string outputExt = "ply";
    int binaryModeVisualizationSteps = 100;
    bool outputRecoveryState = true;
    int outputRecoveryStateStep = 10;
    bool outputRecoveryStateBinary = true;
    bool outputStatistics = false;
    int debugVerboseLvl = 1;
};

class VBDClothPhysicsB200 {
public:
    VBDClothPhysicsB200() = default;
    ~VBDClothPhysicsB200();
    VBDClothPhysicsB200(const VBDClothPhysicsB200&) = delete;
    VBDClothPhysicsB200& operator=(const VBDClothPhysicsB200&) = delete;

    void setRepoRoot(const std::string& root) { repoRoot_ = root; }
    void setDevice(int device) { device_ = device; }

    void loadRunningparameters(std::string inModelInputFile, std::string inParameterFile, std::string outFolder);
    void loadMeshes(int64_t* nVertsTotal, int64_t* nFacesTotal);   // host only: .obj files + colourings (part of initialize)
    void initialize();
    void recoverFromState(const std::string& stateFile);
    void runStep();                       // one frame: numSubsteps x (contact detection, initial guess, colour sweeps, velocities)
    void simulate();                      // writeOutputs(0); while frameId < numFrames: runStep, ++frameId, writeOutputs
    void writeOutputs(const std::string& outFolder, int frameId);
    void saveRecoveryState(const std::string& file, bool binary);
    void writeStatistics(const std::string& outFolder, int frameId);

    gaia_cloth_physics_params& physicsParams() { return phys_; }
    ClothHostRunParams& runParams() { return run_; }
    std::vector<ClothModelInput>& models() { return models_; }
    gaia_cloth_handle* handle() { return h_; }
    int frameId() const { return frameId_; }
    const std::string& outputFolder() const { return outputFolder_; }

private:
    void check(int rc, const char* what);
    std::string resolve(const std::string& p) const;

    std::string repoRoot_, outputFolder_;
    int device_ = 0;
    gaia_cloth_physics_params phys_;
    ClothHostRunParams run_;
    std::vector<ClothModelInput> models_;
    std::vector<HostTriMeshData> meshCache_;
    std::map<std::string, std::vector<std::vector<int32_t>>> colourings_;
    std::vector<ClothDeformerInput> deformers_;
    std::vector<ColliderInput> colliders_;
    ClothAttachmentInput attachments_;
    gaia_cloth_handle* h_ = nullptr;
    int frameId_ = 0;
    double curTime_ = 0.0;
    bool loaded_ = false;
};

void loadObj(const std::string& path, HostTriMeshData& out);

}  // namespace gaia_b200
