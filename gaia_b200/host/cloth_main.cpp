// gaia_vbdcloth_b200 — command-line driver with the reference's VBDCloth calling convention (Simulator/VBDCloth/main.cpp,
// M/Parser/Parser.h:7-45):
//     gaia_vbdcloth_b200 <Models.json> <Parameters.json> <outFolder|noOutput> [-R|--repoRoot DIR] [-r|--recoveryState FILE] [--device N]
// --gui is accepted and ignored (no graphics stack on this box); --CPU is refused (this implementation has no CPU path).
#include <cstdlib>
#include <iostream>
#include <string>
#include <vector>

#include "VBDClothPhysicsB200.h"

int main(int argc, char** argv) {
    std::vector<std::string> pos;
    std::string repoRoot, recoveryStateFile;
    int device = 0, frames = -1;
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
        else if (a == "--frames") frames = std::atoi(next("--frames").c_str());
        else if (a == "-g" || a == "--gui") std::cerr << "[gaia_b200] --gui ignored: the viewer needs a graphics stack, unavailable on this box\n";
        else if (a == "--CPU") { std::cerr << "[gaia_b200] --CPU refused: this implementation of the VBD cloth step has no CPU path\n"; return 2; }
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
        gaia_b200::VBDClothPhysicsB200 physics;
        physics.setRepoRoot(repoRoot);
        physics.setDevice(device);
        physics.loadRunningparameters(pos[0], pos[1], pos[2]);
        if (pos[2] == "noOutput") physics.runParams().saveOutputs = false;
        if (frames >= 0) physics.runParams().numFrames = frames;
        if (dryRun) {
            int64_t nV = 0, nF = 0;
            physics.loadMeshes(&nV, &nF);
            std::cout << "dryRun: " << physics.models().size() << " models, " << nV << " vertices, " << nF << " triangles, numFrames "
                      << physics.runParams().numFrames << ", numSubsteps " << physics.physicsParams().numSubsteps << ", iterations "
                      << physics.physicsParams().iterations << ", contactRadius " << physics.physicsParams().contactRadius << ", outputExt "
                      << physics.runParams().outputExt << std::endl;
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
