// One tet mesh partitioned over the GPUs of one box, driven from ONE process: a host thread per GPU, each holding one slab
// partition (gaia_b200_partition_*) behind its own C-ABI handle; the partitions exchange the positions of their boundary
// vertices per colour through peer memory (gaia_b200_halo_*_pointers). Same inputs, same output files as VBDPhysicsB200 -
// the frames are bit-identical to the single-GPU run. The reference has no counterpart: VBDPhysics owns one
// stream on one device (M/VBD/VBDPhysics.h:212).
#pragma once
#include <array>
#include <functional>

#include "VBDPhysicsB200.h"

namespace gaia_b200 {

class VBDPhysicsB200MultiGPU : public VBDPhysicsB200 {
public:
    explicit VBDPhysicsB200MultiGPU(int nGpus) : nGpus_(nGpus) {}
    ~VBDPhysicsB200MultiGPU() override;
    void initialize();
    void runStep();
    void simulate();
    void writeOutputs(const std::string& outFolder, int frameId) override;
    int numGpus() const { return nGpus_; }
    // --layout replicated: the solve stays partitioned, but every GPU holds the WHOLE mesh as collision geometry and receives
    // every colour's new positions from every other GPU - contacts between any two slabs are seen
    void setReplicatedLayout(bool on) { replicated_ = on; }

protected:
    void fetchSurfaceTopology() override;

private:
    struct Part {
        gaia_handle* h = nullptr;
        gaia_partition* p = nullptr;
        int device = 0;
        std::vector<int32_t> localVerts;     // global ids
        std::vector<uint8_t> owned;
    };
    void forAllParts(const std::function<void(int)>& fn);      // one host thread per partition, joined before returning
    void initializeReplicated(const std::vector<int32_t>& owner, const std::vector<int32_t>& cOffs, const std::vector<int32_t>& cFlat, int nColours);
    int nGpus_ = 1;
    bool replicated_ = false;
    std::vector<Part> parts_;
    std::vector<int32_t> surfIds_;            // global surface vertex ids of the mesh (ascending)
    std::vector<float> global_, surface_;     // assembled positions
};

}  // namespace gaia_b200
