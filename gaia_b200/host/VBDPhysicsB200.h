// VBDPhysicsB200 — C++ host adapter with the reference's VBDPhysics entry points on top of the C ABI of
// libgaia_b200.so (include/gaia_b200.h). A Gaia-style main (S/VBDDynamics/main.cpp:8-34) works unchanged in shape:
//
//     VBDPhysicsB200 physics;
//     physics.loadRunningparameters(modelsJson, parametersJson, outFolder);
//     physics.initialize();
//     if (!recoveryStateFile.empty()) physics.recoverFromState(recoveryStateFile);
//     physics.simulate();
//
// Mirrors (names, argument meaning, file formats):
//   BasePhysicFramework::loadRunningparameters / parseRunningParameters   M/Framework/BasePhysicsFramework.cpp:595-626
//   BasePhysicFramework::initialize / simulate / writeOutputs              M/Framework/BasePhysicsFramework.cpp:101-202, 284-393, 442-506
//   VBDPhysics::runStep / enableModels                                     M/VBD/VBDPhysics.cpp:809-837, 685-709
//   PhysicsState (recovery states, text and binary)                        M/Framework/BasePhysicsFramework.h:127-319
//   writeAllToPLY / writeAllToBinary / saveAsPly                           M/IO/FileIO.cpp:8-185
//   loadDeformers                                                          M/VBD/VBDDeformer.h:179-222
//   ${REPO_ROOT} substitution                                              M/Parser/InputHandler.h:139-183
// The viewer (polyscope / OpenGL) is not part of the path: unavailable on this box.
#pragma once
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "gaia_b200.h"

namespace gaia_b200 {

struct ModelInput {
    std::string materialName = "NeoHookean";
    std::string path, verticesColoringCategoriesPath;
    gaia_object_params obj;
    std::vector<int32_t> fixedPoints;
    // loaded mesh (shared between models with the same path)
    int meshCacheId = -1;
};

struct HostMeshData {
    std::string path;
    std::vector<float> verts;      // 3*nV as read from the .t file
    std::vector<int32_t> tets;     // 4*nT
};

// BasePhysicsParams fields the host loop (not the device step) consumes
struct HostRunParams {
    int numFrames = 5000;
    bool saveOutputs = true;
    // PhysicsParams "applyLocalLineSearch" (an addition of this host: the reference selects it at compile time with
    // APPLY_LOCAL_LINE_SEARCH, M/VBD/VBDPhysics.cpp:25) + the reference's own tau / tau_GD / backtracingLineSearchMaxIters
    bool applyLocalLineSearch = false;
    float tau = 0.5f, tauGD = 0.1f;
    int backtracingLineSearchMaxIters = 16;
    bool saveAllModelsTogether = true;
    std::string outputExt = "ply";
    int binaryModeVisualizationSteps = 100;
    bool outputRecoveryState = true;
    int outputRecoveryStateStep = 10;
    bool outputRecoveryStateBinary = true;
    bool outputStatistics = false;
    int debugVerboseLvl = 1;
};

class VBDPhysicsB200 {
public:
    VBDPhysicsB200() = default;
    virtual ~VBDPhysicsB200();
    VBDPhysicsB200(const VBDPhysicsB200&) = delete;
    VBDPhysicsB200& operator=(const VBDPhysicsB200&) = delete;

    // -R / --repoRoot of the reference CLI; must be set before loadRunningparameters
    void setRepoRoot(const std::string& root) { repoRoot_ = root; }
    void setDevice(int device) { device_ = device; }

    void loadRunningparameters(std::string inModelInputFile, std::string inParameterFile, std::string outFolder);
    void loadMeshes(int64_t* nVertsTotal, int64_t* nTetsTotal);   // host only: .t files + colourings (part of initialize)
    void initialize();
    void recoverFromState(const std::string& stateFile);
    void runStep();                       // one frame: enableModels + numSubsteps x (collision detection, iterations, velocities)
    void simulate();                      // writeOutputs(0); while frameId < numFrames: runStep, ++frameId, writeOutputs
    virtual void writeOutputs(const std::string& outFolder, int frameId);
    void saveRecoveryState(const std::string& file, bool binary);
    void writeStatistics(const std::string& outFolder, int frameId);

    gaia_physics_params& physicsParams() { return phys_; }
    const float* bowlCenter() const { return bowlCenter_; }
    float bowlRadius() const { return bowlRadius_; }
    bool bowlCap() const { return bowlCap_; }
    gaia_collision_params& collisionParams() { return col_; }
    HostRunParams& runParams() { return run_; }
    std::vector<ModelInput>& models() { return models_; }
    gaia_handle* handle() { return h_; }
    int frameId() const { return frameId_; }
    const std::string& outputFolder() const { return outputFolder_; }

protected:
    void check(int rc, const char* what);
    virtual void fetchSurfaceTopology();          // fills surfFaces_ / surfOffset_ (single GPU: from the handle)
    // frame files are written by a second thread from the library's page-locked snapshot ring (two slots): the step of frame
    // f+1 is enqueued while frame f is still being copied and written (FileIO.cpp's writers run inside the frame loop)
    void writeFrameFiles(const std::string& outFolder, int frameId, const float* surfAll);
    void submitFrame(int frameId);
    void writerLoop();
    void finishWriter();
    std::thread writer_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<std::pair<int, int>> jobs_;      // (slot, frame id)
    bool slotBusy_[2] = {false, false};
    bool stopWriter_ = false;
    std::string writerError_;
    std::vector<std::vector<int32_t>> surfFaces_;   // per mesh: surface faces as indices into the surface vertex list (static)
    std::vector<int64_t> surfOffset_;                // per mesh: first surface vertex in the concatenated snapshot
    void loadTetMesh(const std::string& path, HostMeshData& out);
    std::string resolve(const std::string& p) const;

    std::string repoRoot_, outputFolder_;
    int device_ = 0;
    gaia_physics_params phys_;
    float bowlCenter_[3] = {0.f, 0.f, 0.f}, bowlRadius_ = 1.f;      // BasePhysicsParams::bowlCenter / bowlRadius / bowlCap (used with useBowlGround)
    bool bowlCap_ = false;
    gaia_collision_params col_;
    HostRunParams run_;
    std::vector<ModelInput> models_;
    std::vector<HostMeshData> meshCache_;
    std::map<std::string, std::vector<std::vector<int32_t>>> colourings_;
    struct DeformerInput {
        int type = 0;   // 0 Rotator, 1 Translator
        float center[3] = {0, 0, 0}, axis[3] = {0, 0, 0}, speed[3] = {0, 0, 0};
        double angularVelocity = 0, endTime = -1;
        std::vector<int> selectedMeshes;
        std::vector<std::vector<int>> selectedVertices;
    };
    std::vector<DeformerInput> deformers_;
    gaia_handle* h_ = nullptr;
    int frameId_ = 0;
    bool loaded_ = false;
};

// helpers shared with main.cpp and the tests
std::vector<std::vector<int32_t>> loadVertexColoring(const std::string& path);
void saveAsPly(const std::string& filePath, const std::vector<float>& verts, const std::vector<int32_t>& faces);

}  // namespace gaia_b200
