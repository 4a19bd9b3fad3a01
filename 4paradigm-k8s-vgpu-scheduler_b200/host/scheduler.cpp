// scheduler.cpp -- see scheduler.hpp.  Host logic of the Filter path, mirroring the Go code
// it cites; the only call into the GPU library is calcScoreGPU.
#include "scheduler.hpp"

#include <chrono>
#include <unordered_set>
#include <algorithm>
#include <cstring>
#include <ctime>

#include "../../include/vgpuscore.h"

namespace vgpu {

// ---- pkg/util/util.go:110-126 ------------------------------------------------------------
std::string EncodeContainerDevices(const ContainerDevices &cd) {
    std::string tmp;
    for (const ContainerDevice &val : cd)
        tmp += val.UUID + "," + val.Type + "," + std::to_string(val.Usedmem) + "," + std::to_string(val.Usedcores) + ":";
    return tmp;
}
std::string EncodePodDevices(const PodDevices &pd) {
    std::string out;
    for (size_t i = 0; i < pd.size(); i++) {
        if (i) out += ";"; // strings.Join(ss, ";")
        out += EncodeContainerDevices(pd[i]);
    }
    return out;
}

// ---- pkg/util/util.go:128-172, :68-99 (through the library's decoders) --------------------
PodDevices DecodePodDevices(const std::string &str) {
    PodDevices pd;
    const int32_t cap_c = 64, cap_d = 1024;
    std::vector<int32_t> ctr_n(cap_c), uo(cap_d), ul(cap_d), to(cap_d), tl(cap_d), mem(cap_d), cores(cap_d);
    const int32_t nc = vgs_host_decode_pod_devices(str.c_str(), cap_c, cap_d, ctr_n.data(), uo.data(), ul.data(),
                                                   to.data(), tl.data(), mem.data(), cores.data());
    int32_t d = 0;
    for (int32_t c = 0; c < nc; c++) {
        ContainerDevices cd;
        for (int32_t j = 0; j < ctr_n[c]; j++, d++) {
            ContainerDevice x;
            x.UUID = str.substr((size_t)uo[d], (size_t)ul[d]), x.Type = str.substr((size_t)to[d], (size_t)tl[d]);
            x.Usedmem = mem[d], x.Usedcores = cores[d];
            cd.push_back(x);
        }
        pd.push_back(cd);
    }
    return pd;
}
bool DecodeNodeDevices(const std::string &str, std::vector<DeviceInfo> *out) {
    const int32_t cap = 64;
    std::vector<int32_t> cnt(cap), mem(cap), core(cap), numa(cap), health(cap), io(cap), il(cap), to(cap), tl(cap);
    const int32_t n = vgs_host_decode_node_devices(str.c_str(), cap, cnt.data(), mem.data(), core.data(), numa.data(),
                                                   health.data(), io.data(), il.data(), to.data(), tl.data());
    if (n < 0) return false;
    out->clear();
    for (int32_t i = 0; i < n; i++) {
        DeviceInfo d;
        d.ID = str.substr((size_t)io[i], (size_t)il[i]), d.Type = str.substr((size_t)to[i], (size_t)tl[i]);
        d.Index = (uint32_t)i, d.Count = cnt[i], d.Devmem = mem[i], d.Devcore = core[i], d.Numa = numa[i];
        d.Health = health[i] != 0;
        out->push_back(d);
    }
    return true;
}

// ---- pkg/device/nvidia/device.go:114-175 ---------------------------------------------------
static bool lookup(const Container &ctr, const std::string &name, int64_t *v) {
    auto it = ctr.Limits.find(name); // limits take precedence (:119-122)
    if (it == ctr.Limits.end()) {
        it = ctr.Requests.find(name);
        if (it == ctr.Requests.end()) return false;
    }
    *v = it->second;
    return true;
}
ContainerDeviceRequest GenerateResourceRequests(const Container &ctr, const Config &cfg) {
    int64_t n;
    if (!lookup(ctr, cfg.ResourceName, &n)) return ContainerDeviceRequest{};
    int64_t v;
    int memnum = 0;
    if (lookup(ctr, cfg.ResourceMem, &v)) memnum = (int)v;
    int32_t mempnum = 101; // "unset" (:136)
    if (lookup(ctr, cfg.ResourceMemPercentage, &v)) mempnum = (int32_t)v;
    if (mempnum == 101 && memnum == 0) { // :147-153
        if (cfg.DefaultMem != 0) memnum = cfg.DefaultMem;
        else mempnum = 100;
    }
    int32_t corenum = cfg.DefaultCores;
    if (lookup(ctr, cfg.ResourceCores, &v)) corenum = (int32_t)v;
    ContainerDeviceRequest r;
    r.Nums = (int32_t)n;
    r.Type = "NVIDIA";
    r.Memreq = (int32_t)memnum;
    r.MemPercentagereq = mempnum;
    r.Coresreq = corenum;
    return r;
}

// ---- pkg/device/cambricon/device.go:106-137 ---------------------------------------------------
ContainerDeviceRequest GenerateResourceRequestsMLU(const Container &ctr, const Config &cfg) {
    int64_t n;
    if (!lookup(ctr, cfg.MLUResourceCount, &n)) return ContainerDeviceRequest{};
    int64_t v;
    int memnum = 0;
    if (lookup(ctr, cfg.MLUResourceMemory, &v)) memnum = (int)v;
    ContainerDeviceRequest r; // MemPercentagereq and Coresreq stay at Go's zero value
    r.Nums = (int32_t)n;
    r.Type = "MLU";
    r.Memreq = (int32_t)memnum;
    return r;
}

// ---- pkg/device/hygon/device.go:86-135 -----------------------------------------------------------
ContainerDeviceRequest GenerateResourceRequestsDCU(const Container &ctr, const Config &cfg) {
    int64_t n;
    if (!lookup(ctr, cfg.DCUResourceCount, &n)) return ContainerDeviceRequest{};
    int64_t v;
    int memnum = 0;
    if (lookup(ctr, cfg.DCUResourceMemory, &v)) memnum = (int)v;
    int32_t corenum = 0;
    if (lookup(ctr, cfg.DCUResourceCores, &v)) corenum = (int32_t)v;
    ContainerDeviceRequest r;
    r.Nums = (int32_t)n;
    r.Type = "DCU";
    r.Memreq = (int32_t)memnum;
    r.MemPercentagereq = memnum == 0 ? 100 : 0; // :121-124
    r.Coresreq = corenum;
    return r;
}

// ---- pkg/k8sutil/pod.go:26-40 -------------------------------------------------------------------
std::vector<std::vector<ContainerDeviceRequest>> Resourcereqs(const Pod &pod, const Config &cfg) {
    std::vector<std::vector<ContainerDeviceRequest>> counts(pod.Containers.size());
    for (size_t i = 0; i < pod.Containers.size(); i++) {
        const ContainerDeviceRequest vendors[3] = {GenerateResourceRequestsMLU(pod.Containers[i], cfg),
                                                   GenerateResourceRequests(pod.Containers[i], cfg),
                                                   GenerateResourceRequestsDCU(pod.Containers[i], cfg)};
        for (const ContainerDeviceRequest &request : vendors)
            if (request.Nums > 0) counts[i].push_back(request); // :33
    }
    return counts;
}

// ---- registries -----------------------------------------------------------------------------
Scheduler::Scheduler(int device, int max_nodes, bool incremental, const Config &cfg)
    : cfg_(cfg), max_nodes_(max_nodes), incremental_(incremental) {
    vgs_config c;
    memset(&c, 0, sizeof c);
    c.device = device;
    c.max_nodes = max_nodes;
    c.max_devices = VGS_MAX_DEVICES;
    c.max_pods = 1; // Filter scores one pod per call
    c.max_reqs = VGS_MAX_REQS_PER_POD;
    c.max_picks = VGS_MAX_REQS_PER_POD * VGS_MAX_DEVICES;
    if (vgs_init(&c, &ctx_) != VGS_OK) ctx_ = nullptr; // Filter reports it; there is no CPU fallback
}
Scheduler::~Scheduler() { vgs_destroy(ctx_); }

void Scheduler::addNode(const std::string &nodeID, const NodeInfo &info) {
    if (info.Devices.empty()) return; // nodes.go:60-62
    auto it = nodes_.find(nodeID);
    if (it != nodes_.end()) { // :65-70 a known node GAINS the devices (another vendor's register annotation)
        it->second.Devices.insert(it->second.Devices.end(), info.Devices.begin(), info.Devices.end());
    } else {
        nodes_[nodeID] = info;
        nodes_[nodeID].ID = nodeID;
    }
    node_epoch_++;
}

void Scheduler::rmNodeDevice(const std::string &nodeID, const NodeInfo &info) {
    auto it = nodes_.find(nodeID);
    if (it == nodes_.end() || it->second.Devices.empty()) return; // nodes.go:79-83
    std::vector<DeviceInfo> tmp;
    for (const DeviceInfo &val : it->second.Devices) {
        bool found = false;
        for (const DeviceInfo &rm : info.Devices)
            if (val.ID == rm.ID) {
                found = true;
                break;
            }
        if (!found && !val.ID.empty()) tmp.push_back(val); // :94-96
    }
    it->second.Devices = tmp;
    node_epoch_++;
}

NodeInfo Scheduler::registerNode(const std::string &nodeID, const std::string &registerAnnotation) {
    NodeInfo nodeInfo;
    nodeInfo.ID = nodeID;
    std::vector<DeviceInfo> nodedevices;
    if (!DecodeNodeDevices(registerAnnotation, &nodedevices) || nodedevices.empty()) return nodeInfo; // :152-155
    bool found = false; // declared OUTSIDE the device loop in the reference (:195): once one device of the
                        // annotation is already registered, none of the following ones is appended
    for (size_t index = 0; index < nodedevices.size(); index++) {
        const DeviceInfo &deviceinfo = nodedevices[index];
        auto it = nodes_.find(nodeID);
        if (it != nodes_.end()) {
            for (DeviceInfo &val1 : it->second.Devices)
                if (val1.ID == deviceinfo.ID) { // :199-205 refresh the capacities of a known device
                    found = true;
                    if (val1.Devmem != deviceinfo.Devmem || val1.Devcore != deviceinfo.Devcore) node_epoch_++;
                    val1.Devmem = deviceinfo.Devmem;
                    val1.Devcore = deviceinfo.Devcore;
                    break;
                }
        }
        if (!found) {
            DeviceInfo d = deviceinfo;
            d.Index = (uint32_t)index;
            nodeInfo.Devices.push_back(d);
        }
    }
    addNode(nodeID, nodeInfo);
    return nodeInfo;
}

void Scheduler::onAddPod(const Pod &pod, bool terminated) {
    auto node = pod.Annotations.find(AssignedNodeAnnotations);
    if (node == pod.Annotations.end()) return; // scheduler.go:78-81
    auto ids = pod.Annotations.find(AssignedIDsAnnotations);
    if (ids == pod.Annotations.end()) return; // :82-85
    if (terminated) {                         // :86-89 k8sutil.IsPodInTerminatedState
        delPod(pod);
        return;
    }
    addPod(pod, node->second, DecodePodDevices(ids->second)); // :90-91
}
void Scheduler::onDelPod(const Pod &pod) {
    if (!pod.Annotations.count(AssignedNodeAnnotations)) return; // scheduler.go:104-107
    delPod(pod);
}

void Scheduler::queueDelta(const std::string &nodeID, const PodDevices &devs, int sign) {
    if (!incremental_) return;
    auto node = nodes_.find(nodeID);
    if (node == nodes_.end()) return;
    auto row = resident_row_.find(nodeID); // not a row of the resident snapshot: only the usage view follows
    // the resident copy of getNodesUsage's result (what /metrics reads, scheduler.go:232-234) follows the
    // same events, so that a Filter on the resident path does not have to rebuild it
    auto ov = overviewstatus_.find(nodeID);
    for (const ContainerDevices &cd : devs)
        for (const ContainerDevice &u : cd)
            for (size_t s = 0; s < node->second.Devices.size(); s++)
                if (node->second.Devices[s].ID == u.UUID) { // same match as scheduler.go:287-293
                    if (row != resident_row_.end())
                        pending_.push_back({row->second, (int32_t)s, sign, sign * u.Usedmem, sign * u.Usedcores});
                    if (ov != overviewstatus_.end() && s < ov->second.Devices.size()) {
                        DeviceUsage &d = *ov->second.Devices[s];
                        d.Used += sign, d.Usedmem += sign * u.Usedmem, d.Usedcores += sign * u.Usedcores;
                    }
                }
}

void Scheduler::addPod(const Pod &pod, const std::string &nodeID, const PodDevices &devices) {
    if (pods_.count(pod.UID)) return; // pods.go:49-58: only unknown pods are recorded
    pods_[pod.UID] = podInfo{pod.Namespace, pod.Name, pod.UID, nodeID, devices};
    queueDelta(nodeID, devices, +1);
}
void Scheduler::delPod(const Pod &pod) {
    auto it = pods_.find(pod.UID);
    if (it == pods_.end()) return;
    queueDelta(it->second.NodeID, it->second.Devices, -1);
    pods_.erase(it);
}

// ---- scheduler.go:249-310 ---------------------------------------------------------------------
void Scheduler::getNodesUsage(const std::vector<std::string> &nodes, std::map<std::string, NodeUsage *> *cache,
                              std::map<std::string, std::string> *failedNodes) {
    overviewstatus_.clear();
    for (const auto &kv : nodes_) { // :259-278 one fresh DeviceUsage per registered device
        NodeUsage nu;
        for (const DeviceInfo &d : kv.second.Devices) {
            auto u = std::make_shared<DeviceUsage>();
            u->Id = d.ID, u->Index = d.Index, u->Count = d.Count, u->Totalmem = d.Devmem;
            u->Totalcore = d.Devcore, u->Type = d.Type, u->Numa = d.Numa, u->Health = d.Health;
            nu.Devices.push_back(u);
        }
        overviewstatus_[kv.first] = nu;
    }
    for (const auto &kv : pods_) { // :280-297 fold the scheduled pods' usage in, matching by UUID
        auto node = overviewstatus_.find(kv.second.NodeID);
        if (node == overviewstatus_.end()) continue;
        for (const ContainerDevices &ds : kv.second.Devices)
            for (const ContainerDevice &udevice : ds)
                for (auto &d : node->second.Devices)
                    if (d->Id == udevice.UUID) {
                        d->Used++;
                        d->Usedmem += udevice.Usedmem;
                        d->Usedcores += udevice.Usedcores;
                    }
    }
    for (const std::string &nodeID : nodes) { // :299-307
        auto it = overviewstatus_.find(nodeID);
        if (it == overviewstatus_.end()) {
            (*failedNodes)[nodeID] = "node unregisterd";
            continue;
        }
        (*cache)[nodeID] = &it->second;
    }
}

bool Scheduler::uploadSnapshot(const std::vector<std::string> &order, const std::map<std::string, NodeUsage *> &nodes,
                               std::string *err) {
    size_t dmax = 1;
    for (const std::string &id : order) dmax = std::max(dmax, nodes.at(id)->Devices.size());
    if (dmax > VGS_MAX_DEVICES) {
        *err = vgs_strerror(VGS_EDMAX);
        return false;
    }
    const size_t n = order.size();
    std::vector<int32_t> planes[VGS_NUM_PLANES];
    for (auto &p : planes) p.assign(std::max<size_t>(n * dmax, 1), 0);
    std::vector<int32_t> ndev(std::max<size_t>(n, 1), 0);
    resident_row_.clear();
    for (size_t row = 0; row < n; row++) {
        const DeviceUsageList &devs = nodes.at(order[row])->Devices;
        ndev[row] = (int32_t)devs.size();
        resident_row_[order[row]] = (int32_t)row;
        for (size_t s = 0; s < devs.size(); s++) { // registration order == slot
            const DeviceUsage &d = *devs[s];
            const size_t i = row * dmax + s;
            planes[VGS_PLANE_USED][i] = d.Used, planes[VGS_PLANE_COUNT][i] = d.Count;
            planes[VGS_PLANE_USEDMEM][i] = d.Usedmem, planes[VGS_PLANE_TOTALMEM][i] = d.Totalmem;
            planes[VGS_PLANE_USEDCORES][i] = d.Usedcores, planes[VGS_PLANE_TOTALCORE][i] = d.Totalcore;
            planes[VGS_PLANE_NUMA][i] = d.Numa;
            auto t = type_id_.find(d.Type);
            if (t == type_id_.end()) {
                if (types_.size() >= VGS_MAX_TYPES) {
                    *err = "more than VGS_MAX_TYPES distinct device types";
                    return false;
                }
                t = type_id_.emplace(d.Type, (int32_t)types_.size()).first;
                types_.push_back(d.Type);
            }
            planes[VGS_PLANE_TYPEID][i] = t->second;
        }
    }
    const int32_t *ptrs[VGS_NUM_PLANES];
    for (int f = 0; f < VGS_NUM_PLANES; f++) ptrs[f] = planes[f].data();
    std::vector<uint32_t> tflags;
    for (const std::string &t : types_) tflags.push_back(vgs_host_type_flags(t.c_str()));
    int rc = vgs_snapshot_type_flags(ctx_, (int32_t)tflags.size(), tflags.data());
    if (rc == VGS_OK) rc = vgs_snapshot_upload(ctx_, (int32_t)n, (int32_t)dmax, ptrs, ndev.data(), nullptr);
    if (rc != VGS_OK) {
        *err = std::string(vgs_strerror(rc)) + ": " + vgs_last_error(ctx_);
        return false;
    }
    resident_order_ = order;
    resident_epoch_ = node_epoch_;
    pending_.clear();
    return true;
}

// ---- calcScore (score.go:183-214) + pick (scheduler.go:386-387), on the GPU -----------------------
static double now_us() {
    return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

bool Scheduler::calcScoreGPU(const std::vector<std::string> &order, const std::map<std::string, NodeUsage *> *nodes,
                             const std::vector<std::vector<ContainerDeviceRequest>> &nums,
                             const std::map<std::string, std::string> &annos, NodeScore *winner, bool *found,
                             std::string *err) {
    *found = false;
    if (!ctx_) {
        *err = "libvgpuscore: no usable GPU (vgs_init failed); there is no CPU fallback";
        return false;
    }
    if ((int)order.size() > max_nodes_) {
        *err = vgs_strerror(VGS_ETOOBIG);
        return false;
    }
    // snapshot: resident + deltas when nothing but pod usage changed, else a full upload
    double t_gpu = 0.0, t0;
    if (!nodes || (incremental_ && resident_epoch_ == node_epoch_ && resident_order_ == order)) {
        if (!pending_.empty()) {
            t0 = now_us();
            std::vector<vgs_delta> d(pending_.size());
            for (size_t i = 0; i < pending_.size(); i++)
                d[i] = vgs_delta{pending_[i].row, pending_[i].slot, pending_[i].used, pending_[i].mem, pending_[i].cores};
            int rc = vgs_snapshot_delta(ctx_, (int32_t)d.size(), d.data());
            t_gpu += now_us() - t0;
            if (rc != VGS_OK) {
                *err = std::string(vgs_strerror(rc)) + ": " + vgs_last_error(ctx_);
                return false;
            }
            pending_.clear();
        }
    } else {
        t0 = now_us();
        const bool ok = uploadSnapshot(order, *nodes, err);
        t_gpu += now_us() - t0; // (includes packing the planes)
        if (!ok) return false;
    }
    // requests: the string policy (checkType, score.go:71-84) evaluated once per distinct type
    auto anno = [&](const char *k) -> const char * {
        auto it = annos.find(k);
        return it == annos.end() ? nullptr : it->second.c_str();
    };
    std::vector<const char *> type_c;
    for (const std::string &t : types_) type_c.push_back(t.c_str());
    std::vector<vgs_req> reqs;
    int total = 0;
    for (size_t ctr = 0; ctr < nums.size(); ctr++)
        for (const ContainerDeviceRequest &k : nums[ctr]) {
            vgs_req r;
            memset(&r, 0, sizeof r);
            r.ctr = (int32_t)ctr, r.nums = k.Nums, r.memreq = k.Memreq, r.mempct = k.MemPercentagereq;
            r.coresreq = k.Coresreq;
            if (k.Type == "MLU") {
                r.typemask = vgs_host_typemask_mlu(k.Memreq, anno("cambricon.com/use-mlutype"),
                                                   anno("cambricon.com/nouse-mlutype"), type_c.data(),
                                                   (int32_t)type_c.size(), &r.flags);
            } else if (k.Type == "DCU") {
                r.typemask = vgs_host_typemask("DCU", anno("hygon.com/use-dcutype"), anno("hygon.com/nouse-dcutype"),
                                               type_c.data(), (int32_t)type_c.size());
            } else {
                r.typemask = vgs_host_typemask(k.Type.c_str(), anno("nvidia.com/use-gputype"),
                                               anno("nvidia.com/nouse-gputype"), type_c.data(), (int32_t)type_c.size());
            }
            if (k.Type == "NVIDIA" && vgs_host_numa_bind(anno("nvidia.com/numa-bind"))) r.flags |= VGS_REQ_NUMA_BIND;
            reqs.push_back(r);
            total += k.Nums;
        }
    vgs_pod pod{0, (int32_t)reqs.size()};
    vgs_result res;
    std::vector<vgs_pick> picks(std::max(total, 1));
    t0 = now_us();
    int rc = vgs_score_batch(ctx_, 1, &pod, (int32_t)reqs.size(), reqs.data(), &res, picks.data());
    timing_.gpu_us = t_gpu + (now_us() - t0);
    if (rc != VGS_OK) {
        *err = std::string(vgs_strerror(rc)) + ": " + vgs_last_error(ctx_);
        return false;
    }
    if (res.best_row < 0) return true; // len(*nodeScores) == 0
    *found = true;
    winner->nodeID = order[res.best_row];
    winner->score = res.score;
    winner->devices.clear();
    const std::vector<DeviceInfo> &devs = nodes_.at(winner->nodeID).Devices; // registration order == slot
    size_t i = 0;
    for (size_t ctr = 0; ctr < nums.size(); ctr++) { // util.PodDevices: one list per container (score.go:196,201)
        ContainerDevices cd;
        for (const ContainerDeviceRequest &k : nums[ctr])
            for (int j = 0; j < k.Nums; j++, i++) {
                ContainerDevice c;
                c.UUID = devs[picks[i].slot].ID, c.Type = k.Type;
                c.Usedmem = picks[i].usedmem, c.Usedcores = picks[i].usedcores;
                cd.push_back(c);
            }
        winner->devices.push_back(cd);
    }
    return true;
}

// ---- scheduler.go:354-402 ---------------------------------------------------------------------
bool Scheduler::Filter(const ExtenderArgs &args, ExtenderFilterResult *res, std::string *err) {
    *res = ExtenderFilterResult();
    auto nums = Resourcereqs(args.pod, cfg_);
    int total = 0;
    for (const auto &n : nums)
        for (const auto &k : n) total += k.Nums;
    if (total == 0) { // :357-370 nothing of ours requested: every node passes
        res->NodeNames = args.NodeNames;
        return true;
    }
    const double t_begin = now_us();
    timing_ = FilterTiming();
    delPod(args.pod); // :372
    NodeScore m;
    bool found = false;
    // Resident path: the node registry has not changed since the upload and kube-scheduler
    // sends the same candidate list -> the packed snapshot on the GPU, its row order, the per-node usage view
    // and the "unregistered" entries are all still valid; pod events since then are queued deltas.  Nothing of
    // getNodesUsage (scheduler.go:249-310: N x D allocations + a pods x devices UUID scan) is redone.
    const bool fast = incremental_ && ctx_ && resident_epoch_ == node_epoch_ && args.NodeNames == resident_names_;
    timing_.fast = fast;
    bool ok;
    if (fast) {
        res->FailedNodes = resident_failed_;
        ok = calcScoreGPU(resident_order_, nullptr, nums, args.pod.Annotations, &m, &found, err);
    } else {
        std::map<std::string, NodeUsage *> nodeUsage;
        getNodesUsage(args.NodeNames, &nodeUsage, &res->FailedNodes);
        std::vector<std::string> order; // rows = args.NodeNames order, registered nodes only, each once
        std::unordered_set<std::string> seen;
        order.reserve(args.NodeNames.size());
        for (const std::string &id : args.NodeNames)
            if (nodeUsage.count(id) && seen.insert(id).second) order.push_back(id);
        ok = calcScoreGPU(order, &nodeUsage, nums, args.pod.Annotations, &m, &found, err);
        if (ok && incremental_) {
            resident_names_ = args.NodeNames;
            resident_failed_ = res->FailedNodes;
        }
    }
    timing_.total_us = now_us() - t_begin;
    if (!ok) return false; // :378-380
    if (!found) return true; // :381-385 ExtenderFilterResult{FailedNodes}
    std::map<std::string, std::string> &a = res->PatchedAnnotations; // :388-393
    a[AssignedNodeAnnotations] = m.nodeID;
    a[AssignedTimeAnnotations] = std::to_string((long long)time(nullptr));
    a[AssignedIDsAnnotations] = EncodePodDevices(m.devices);
    a[AssignedIDsToAllocateAnnotations] = a[AssignedIDsAnnotations];
    addPod(args.pod, m.nodeID, m.devices); // :394
    res->FailedNodes.clear();              // the success result carries NodeNames only (:400)
    res->NodeNames = {m.nodeID};
    return true;
}

} // namespace vgpu

// ---- flat C façade so that the Python tests can drive the C++ host ----------------------------
extern "C" {

void *vgh_new(int device, int max_nodes, int incremental, int default_mem, int default_cores) {
    vgpu::Config cfg;
    cfg.DefaultMem = default_mem;
    cfg.DefaultCores = default_cores;
    return new vgpu::Scheduler(device, max_nodes, incremental != 0, cfg);
}
void vgh_free(void *s) { delete static_cast<vgpu::Scheduler *>(s); }

void vgh_add_node(void *s, const char *node_id, int n, const char *const *ids, const int32_t *count,
                  const int32_t *devmem, const int32_t *devcore, const char *const *types, const int32_t *numa) {
    vgpu::NodeInfo info;
    for (int i = 0; i < n; i++) {
        vgpu::DeviceInfo d;
        d.ID = ids[i], d.Index = (uint32_t)i, d.Count = count[i], d.Devmem = devmem[i], d.Devcore = devcore[i];
        d.Type = types[i], d.Numa = numa[i];
        info.Devices.push_back(d);
    }
    static_cast<vgpu::Scheduler *>(s)->addNode(node_id, info);
}

void vgh_last_timing(void *s, double *total_us, double *gpu_us, int *fast) {
    const vgpu::Scheduler::FilterTiming &t = static_cast<vgpu::Scheduler *>(s)->lastTiming();
    *total_us = t.total_us, *gpu_us = t.gpu_us, *fast = t.fast ? 1 : 0;
}

void vgh_del_pod(void *s, const char *uid) {
    vgpu::Pod p;
    p.UID = uid;
    static_cast<vgpu::Scheduler *>(s)->delPod(p);
}

// one vendor's register annotation of one node; returns how many devices were new
int vgh_register_node(void *s, const char *node_id, const char *anno) {
    return (int)static_cast<vgpu::Scheduler *>(s)->registerNode(node_id, anno).Devices.size();
}
// drop the devices named in a register annotation (a vendor's devices leaving, scheduler.go:166)
void vgh_rm_node_devices(void *s, const char *node_id, const char *anno) {
    vgpu::NodeInfo info;
    if (!vgpu::DecodeNodeDevices(anno, &info.Devices)) return;
    static_cast<vgpu::Scheduler *>(s)->rmNodeDevice(node_id, info);
}
// pod informer event: node_anno / ids_anno NULL = annotation absent
void vgh_on_add_pod(void *s, const char *ns, const char *name, const char *uid, const char *node_anno,
                    const char *ids_anno, int terminated) {
    vgpu::Pod pod;
    pod.Namespace = ns, pod.Name = name, pod.UID = uid;
    if (node_anno) pod.Annotations[vgpu::AssignedNodeAnnotations] = node_anno;
    if (ids_anno) pod.Annotations[vgpu::AssignedIDsAnnotations] = ids_anno;
    static_cast<vgpu::Scheduler *>(s)->onAddPod(pod, terminated != 0);
}
// res: [n_ctrs][9] container limits (order below), INT64_MIN = absent.
// out: "node=<id>\nids=<vgpu-ids-new>\nfailed=<id>:<why>;...\npass=<id>,<id>...\n" ; returns 0, or -1 with
// out = error text (the reference's `return nil, err`).
int vgh_filter(void *s, const char *ns, const char *name, const char *uid, int n_annos, const char *const *keys,
               const char *const *vals, int n_ctrs, const int64_t *res, int n_nodes, const char *const *node_names,
               char *out, int cap) {
    vgpu::ExtenderArgs args;
    args.pod.Namespace = ns, args.pod.Name = name, args.pod.UID = uid;
    for (int i = 0; i < n_annos; i++) args.pod.Annotations[keys[i]] = vals[i];
    // res[c][9]: the container's limits in this order, INT64_MIN = absent
    static const char *rn[9] = {"nvidia.com/gpu",       "nvidia.com/gpumem",    "nvidia.com/gpumem-percentage",
                                "nvidia.com/gpucores",  "cambricon.com/mlunum", "cambricon.com/mlumem",
                                "hygon.com/dcunum",     "hygon.com/dcumem",     "hygon.com/dcucores"};
    for (int c = 0; c < n_ctrs; c++) {
        vgpu::Container ctr;
        for (int k = 0; k < 9; k++)
            if (res[c * 9 + k] != INT64_MIN) ctr.Limits[rn[k]] = res[c * 9 + k];
        args.pod.Containers.push_back(ctr);
    }
    for (int i = 0; i < n_nodes; i++) args.NodeNames.push_back(node_names[i]);
    vgpu::ExtenderFilterResult r;
    std::string err, text;
    if (!static_cast<vgpu::Scheduler *>(s)->Filter(args, &r, &err)) {
        snprintf(out, (size_t)cap, "%s", err.c_str());
        return -1;
    }
    if (!r.PatchedAnnotations.empty()) {
        text += "node=" + r.PatchedAnnotations[vgpu::AssignedNodeAnnotations] + "\n";
        text += "ids=" + r.PatchedAnnotations[vgpu::AssignedIDsAnnotations] + "\n";
    }
    text += "failed=";
    for (const auto &kv : r.FailedNodes) text += kv.first + ":" + kv.second + ";";
    text += "\npass=";
    for (size_t i = 0; i < r.NodeNames.size(); i++) text += (i ? "," : "") + r.NodeNames[i];
    text += "\n";
    if ((int)text.size() + 1 > cap) return -2;
    memcpy(out, text.c_str(), text.size() + 1);
    return 0;
}

} // extern "C"
