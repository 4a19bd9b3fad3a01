// scheduler.hpp -- C++ mirror of the extender's Filter path, above the C ABI.
//
// The reference host code is Go (pkg/scheduler); no Go toolchain exists on this image, so
// this is the same host logic in C++ with the reference's names, argument meaning and error
// behaviour.  Only calcScore + the best-node pick run on the GPU (libvgpuscore.so); everything
// else here is what stays on the host in the Go extender too.
//
//   types            pkg/util/types.go:85-117, pkg/scheduler/nodes.go:28-48, score.go:29-33
//   Resourcereqs     pkg/k8sutil/pod.go:26-40 + the vendors' GenerateResourceRequests (nvidia/device.go:114-175,
//                    cambricon/device.go:106-137, hygon/device.go:86-135)
//   getNodesUsage    pkg/scheduler/scheduler.go:249-310
//   Filter           pkg/scheduler/scheduler.go:354-402
//   EncodePodDevices pkg/util/util.go:110-126
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

struct vgs_ctx;

namespace vgpu {

// ---- pkg/util/types.go ---------------------------------------------------------------
struct ContainerDevice { // :85-91
    int Idx = 0;
    std::string UUID, Type;
    int32_t Usedmem = 0, Usedcores = 0;
};
struct ContainerDeviceRequest { // :93-99
    int32_t Nums = 0;
    std::string Type;
    int32_t Memreq = 0, MemPercentagereq = 0, Coresreq = 0;
};
using ContainerDevices = std::vector<ContainerDevice>;
using PodDevices = std::vector<ContainerDevices>;
struct DeviceUsage { // :105-117
    std::string Id;
    uint32_t Index = 0;
    int32_t Used = 0, Count = 0, Usedmem = 0, Totalmem = 0, Totalcore = 0, Usedcores = 0;
    int Numa = 0;
    std::string Type;
    bool Health = true;
};

// ---- pkg/scheduler/nodes.go, pods.go ---------------------------------------------------
struct DeviceInfo { // nodes.go:28-37
    std::string ID;
    uint32_t Index = 0;
    int32_t Count = 0, Devmem = 0, Devcore = 0;
    std::string Type;
    int Numa = 0;
    bool Health = true;
};
struct NodeInfo { // nodes.go:39-42
    std::string ID;
    std::vector<DeviceInfo> Devices;
};
using DeviceUsageList = std::vector<std::shared_ptr<DeviceUsage>>; // nodes.go:44 (slice of pointers)
struct NodeUsage {                                                 // nodes.go:46-48
    DeviceUsageList Devices;
};
struct podInfo { // pods.go:28-35
    std::string Namespace, Name, UID, NodeID;
    PodDevices Devices;
};
struct NodeScore { // score.go:29-33
    std::string nodeID;
    PodDevices devices;
    float score = 0;
};

// ---- the slice of the k8s API objects Filter reads ----------------------------------------
struct Container { // corev1.Container.Resources, Limits before Requests (nvidia/device.go:119-122)
    std::map<std::string, int64_t> Limits, Requests;
};
struct Pod {
    std::string Namespace, Name, UID;
    std::map<std::string, std::string> Annotations;
    std::vector<Container> Containers;
};
struct ExtenderArgs { // extenderv1.ExtenderArgs with nodeCacheCapable: only NodeNames is set
    Pod pod;
    std::vector<std::string> NodeNames;
};
struct ExtenderFilterResult { // extenderv1.ExtenderFilterResult
    std::vector<std::string> NodeNames;
    std::map<std::string, std::string> FailedNodes;
    std::string Error;
    // what Filter patches onto the pod (scheduler.go:388-395), returned instead of an API call
    std::map<std::string, std::string> PatchedAnnotations;
};

// pkg/scheduler/config/config.go:19-24 and the vendor flag defaults (nvidia/device.go:41-47)
struct Config {
    int32_t DefaultMem = 0, DefaultCores = 0;
    std::string ResourceName = "nvidia.com/gpu", ResourceMem = "nvidia.com/gpumem",
                ResourceMemPercentage = "nvidia.com/gpumem-percentage", ResourceCores = "nvidia.com/gpucores";
    // cambricon/device.go:40-43, hygon/device.go:34-38
    std::string MLUResourceCount = "cambricon.com/mlunum", MLUResourceMemory = "cambricon.com/mlumem";
    std::string DCUResourceCount = "hygon.com/dcunum", DCUResourceMemory = "hygon.com/dcumem",
                DCUResourceCores = "hygon.com/dcucores";
};

// annotation keys, pkg/util/types.go:26-31
constexpr const char *AssignedTimeAnnotations = "4pd.io/vgpu-time";
constexpr const char *AssignedIDsAnnotations = "4pd.io/vgpu-ids-new";
constexpr const char *AssignedIDsToAllocateAnnotations = "4pd.io/devices-to-allocate";
constexpr const char *AssignedNodeAnnotations = "4pd.io/vgpu-node";

std::string EncodeContainerDevices(const ContainerDevices &cd); // util.go:110-118
std::string EncodePodDevices(const PodDevices &pd);             // util.go:120-126
PodDevices DecodePodDevices(const std::string &str);           // util.go:158-172 (errors come back as an empty list)
// util.go:68-99; false for what the reference rejects
bool DecodeNodeDevices(const std::string &str, std::vector<DeviceInfo> *out);

// nvidia/device.go:114-175; a zero-valued request (Nums == 0) means "this vendor is not asked for"
ContainerDeviceRequest GenerateResourceRequests(const Container &ctr, const Config &cfg);
ContainerDeviceRequest GenerateResourceRequestsMLU(const Container &ctr, const Config &cfg); // cambricon/device.go:106-137
ContainerDeviceRequest GenerateResourceRequestsDCU(const Container &ctr, const Config &cfg); // hygon/device.go:86-135
// k8sutil/pod.go:26-40.  The reference walks the Go map device.GetDevices() (devices.go:43-52), whose
// order is random per call; here the vendors come in the fixed order Cambricon, Nvidia, Hygon.
std::vector<std::vector<ContainerDeviceRequest>> Resourcereqs(const Pod &pod, const Config &cfg);

class Scheduler {
  public:
    // incremental: keep the packed snapshot resident on the GPU and fold addPod/delPod in as
    // usage deltas (vgs_snapshot_delta) instead of re-uploading it on every Filter
    Scheduler(int device, int max_nodes, bool incremental = false, const Config &cfg = Config());
    ~Scheduler();
    Scheduler(const Scheduler &) = delete;
    Scheduler &operator=(const Scheduler &) = delete;

    void addNode(const std::string &nodeID, const NodeInfo &info);      // nodes.go:59-74
    void rmNodeDevice(const std::string &nodeID, const NodeInfo &info); // nodes.go:76-101
    // the data half of RegisterFromNodeAnnotatons for one node and one vendor's register annotation
    // (scheduler.go:152-155, 192-222; the handshake patches :157-191 are control plane).  Returns the
    // NodeInfo handed to addNode (the devices that were new).
    NodeInfo registerNode(const std::string &nodeID, const std::string &registerAnnotation);
    // pod informer callbacks (scheduler.go:72-108): the usage source of getNodesUsage
    void onAddPod(const Pod &pod, bool terminated);
    void onDelPod(const Pod &pod);
    void addPod(const Pod &pod, const std::string &nodeID, const PodDevices &devices); // pods.go:46-60
    void delPod(const Pod &pod);                                                        // pods.go:62-70

    // scheduler.go:354-402.  Returns false + err when the reference would return (nil, err).
    bool Filter(const ExtenderArgs &args, ExtenderFilterResult *res, std::string *err);

    const std::map<std::string, NodeUsage> &InspectAllNodesUsage() const { return overviewstatus_; } // :232-234
    // wall time of the last Filter: the whole call, and the part spent inside libvgpuscore calls (microseconds);
    // fast = it took the resident path (no getNodesUsage rebuild, no snapshot upload)
    struct FilterTiming { double total_us = 0, gpu_us = 0; bool fast = false; };
    const FilterTiming &lastTiming() const { return timing_; }

  private:
    // scheduler.go:249-310
    void getNodesUsage(const std::vector<std::string> &nodes, std::map<std::string, NodeUsage *> *cache,
                       std::map<std::string, std::string> *failedNodes);
    // score.go:183-214 + scheduler.go:386-387 on the GPU: winner or nullptr-equivalent (found = false)
    // (nodes == nullptr: the resident path -- snapshot and row order are the ones of the last upload)
    bool calcScoreGPU(const std::vector<std::string> &order, const std::map<std::string, NodeUsage *> *nodes,
                      const std::vector<std::vector<ContainerDeviceRequest>> &nums,
                      const std::map<std::string, std::string> &annos, NodeScore *winner, bool *found,
                      std::string *err);
    bool uploadSnapshot(const std::vector<std::string> &order, const std::map<std::string, NodeUsage *> &nodes,
                        std::string *err);

    Config cfg_;
    vgs_ctx *ctx_ = nullptr;
    int max_nodes_;
    bool incremental_;
    std::map<std::string, NodeInfo> nodes_;  // nodeManager.nodes
    std::map<std::string, podInfo> pods_;    // podManager.pods, keyed by UID
    std::map<std::string, NodeUsage> overviewstatus_;
    // packed-snapshot bookkeeping
    std::vector<std::string> types_;               // distinct DeviceUsage.Type strings
    std::map<std::string, int32_t> type_id_;
    std::vector<std::string> resident_order_;      // rows of the resident snapshot
    std::unordered_map<std::string, int32_t> resident_row_;
    std::vector<std::string> resident_names_;      // args.NodeNames of the Filter that uploaded it, verbatim
    std::map<std::string, std::string> resident_failed_; // ... and its "node unregisterd" entries
    FilterTiming timing_;
    uint64_t node_epoch_ = 0, resident_epoch_ = ~0ull;
    struct Delta { int32_t row, slot, used, mem, cores; };
    std::vector<Delta> pending_;                   // usage changes since the last upload
    void queueDelta(const std::string &nodeID, const PodDevices &devs, int sign);
};

} // namespace vgpu
