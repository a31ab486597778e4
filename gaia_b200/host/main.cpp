// gaia_vbd_b200 — command-line driver with the reference's VBDDynamics calling convention (S/VBDDynamics/main.cpp:8-34,
// M/Parser/Parser.h:7-45):
//     gaia_vbd_b200 <Models.json> <Parameters.json> <outFolder|noOutput> [-R|--repoRoot DIR] [-r|--recoveryState FILE] [--device N] [--gpus N [--layout ring|replicated]]
// --gpus N (N > 1): the scene's single tet mesh is cut into N slabs, one per GPU of this box (VBDPhysicsB200MultiGPU);
// --layout replicated: every GPU also holds the whole mesh as collision geometry, so contacts between any two slabs are seen.
// --gui is accepted and ignored (no graphics stack on this box); --CPU is refused (this implementation has no CPU path).
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "VBDPhysicsB200MultiGPU.h"

int main(int argc, char** argv) {
    std::vector<std::string> pos;
    std::string repoRoot, recoveryStateFile;
    int device = 0, gpus = 1;
    std::string layout = "ring";
    bool dryRun = false;   // parse the inputs and load the meshes, print the scene size, do not touch the GPU
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&](const char* what) -> std::string {
            if (i + 1 >= argc) { std::cerr << "missing value for " << what << "\n"; std::exit(1); }
            return argv[++i];
        };
        if (a == "-R" || a == "--repoRoot") repoRoot = next("--repoRoot");
        else if (a == "-r" || a == "--recoveryState") recoveryStateFile = next("--recoveryState");
        else if (a == "--device") device = std::atoi(next("--device").c_str());
        else if (a == "--gpus") gpus = std::atoi(next("--gpus").c_str());
        else if (a == "--layout") { layout = next("--layout"); if (layout != "ring" && layout != "replicated") { std::cerr << "--layout ring | replicated\n"; return 1; } }
        else if (a == "-g" || a == "--gui") std::cerr << "[gaia_b200] --gui ignored: the viewer needs a graphics stack, unavailable on this box\n";
        else if (a == "--CPU") { std::cerr << "[gaia_b200] --CPU refused: this implementation of the VBD step has no CPU path\n"; return 2; }
        else if (a == "--git") {}
        else if (a == "--dryRun") dryRun = true;
        else if (!a.empty() && a[0] == '-') { std::cerr << "error parsing options: unknown option " << a << "\n"; return 1; }
        else pos.push_back(a);
    }
    if (pos.size() < 3) {
        std::cout << "[Error] Need at least 3 input parameters!" << std::endl;
        return -1;
    }
    try {
        if (gpus > 1) {
            if (!recoveryStateFile.empty()) { std::cerr << "[gaia_b200] --recoveryState is not supported together with --gpus\n"; return 2; }
            gaia_b200::VBDPhysicsB200MultiGPU physics(gpus);
            physics.setReplicatedLayout(layout == "replicated");
            physics.setRepoRoot(repoRoot);
            physics.loadRunningparameters(pos[0], pos[1], pos[2]);
            if (pos[2] == "noOutput") physics.runParams().saveOutputs = false;
            physics.initialize();
            std::cout << "----------------------------------------------------\nOutput folder is: " << physics.outputFolder() << " (" << gpus << " GPUs)" << std::endl;
            physics.simulate();
            return 0;
        }
        gaia_b200::VBDPhysicsB200 physics;
        physics.setRepoRoot(repoRoot);
        physics.setDevice(device);
        physics.loadRunningparameters(pos[0], pos[1], pos[2]);
        if (pos[2] == "noOutput") physics.runParams().saveOutputs = false;
        if (dryRun) {
            int64_t nV = 0, nT = 0;
            physics.loadMeshes(&nV, &nT);
            std::cout << "dryRun: " << physics.models().size() << " models, " << nV << " vertices, " << nT << " tets, numFrames "
                      << physics.runParams().numFrames << ", numSubsteps " << physics.physicsParams().numSubsteps << ", iterations "
                      << physics.physicsParams().iterations << ", outputExt " << physics.runParams().outputExt << std::endl;
            if (physics.physicsParams().useBowlGround)
                std::cout << "dryRun: bowl ground, centre (" << physics.bowlCenter()[0] << ", " << physics.bowlCenter()[1] << ", " << physics.bowlCenter()[2]
                          << "), radius " << physics.bowlRadius() << ", cap " << (physics.bowlCap() ? 1 : 0) << std::endl;
            return 0;
        }
        physics.initialize();
        if (!recoveryStateFile.empty()) physics.recoverFromState(recoveryStateFile);
        std::cout << "----------------------------------------------------\nOutput folder is: " << physics.outputFolder() << std::endl;
        physics.simulate();
    } catch (const std::exception& e) {
        std::cerr << "[gaia_b200] fatal: " << e.what() << std::endl;
        return 3;
    }
    return 0;
}
