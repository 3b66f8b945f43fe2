// dealer_host.h — host-side mirror of nano-gpu-scheduler's `dealer.Dealer` interface in C++.
//
// The reference's host language is Go (pkg/dealer/dealer.go:23-43); no Go toolchain exists in
// this image, so the layer a maintainer would write in Go is mirrored here in
// C++ with the same method names, argument meaning and error behaviour, on top of the C ABI of
// include/ngs_dealer.h.  Everything numeric is a C-ABI call; this file only keeps what the Go
// DealerImpl keeps: node name -> row, PodMaps / ReleasedPodMap, the usage strings, error texts.
// Kubernetes objects are reduced to the fields the dealer reads.
//
// Citations are relative to /root/reference/pkg/dealer/ unless stated.
#pragma once
#include <cstdint>
#include <map>
#include <set>
#include <string>
#include <utility>
#include <vector>

#include "../../include/ngs_dealer.h"

namespace ngs_host {

// v1.Container / v1.Pod / v1.Node, reduced
struct Container {
    std::string name;
    bool has_gpu_limit = false;  // Resources.Limits["nano-gpu/gpu-percent"] present (pkg/utils/pod.go:94-100)
    int64_t gpu_percent = 0;
};
struct Pod {
    std::string ns, name, uid, node_name;  // node_name = Spec.NodeName
    std::vector<Container> containers;
    std::map<std::string, std::string> annotations, labels;
};
struct Node {
    std::string name;
    bool has_capacity = false;  // Status.Capacity["nano-gpu/gpu-percent"] present
    int64_t gpu_percent_capacity = 0;
};

// PolicySpec.SyncPeriod (type.go:16-27)
struct Period {
    std::string name;
    int64_t period_ns = 0;
};
struct PolicySpec {
    std::vector<Period> sync_period;
};

// what NodeLister.Get and the apiserver pod list give getNodeInfo (dealer.go:271-301)
class ClusterSource {
  public:
    virtual ~ClusterSource() {}
    virtual bool GetNode(const std::string& name, Node* out) = 0;                             // false: `node "x" not found`
    virtual std::vector<Pod> ListAssumedPods(const std::string& node_name) { return {}; }    // label nano-gpu/assume=true
};

enum class RaterKind { Binpack, Spread, Sample };  // controller.Rater (pkg/controller/controller.go:38)

struct GPUCoreUsage {
    std::string CoreUsage, UpdateTime;
};  // nodeusage.go:10-13
struct GPUMemoryUsage {
    std::string MemoryUsage, UpdateTime;
};  // nodeusage.go:15-18

// An error is a string; "" is Go's nil.
using Error = std::string;

class Dealer {
  public:
    // NewDealer (dealer.go:45-75)
    Dealer(ClusterSource* src, RaterKind rater, uint32_t max_nodes, int device = 0);
    ~Dealer();
    bool ok() const { return h_ != nullptr; }
    const std::string& init_error() const { return init_error_; }

    // the 19 methods of dealer.Dealer, same order as dealer.go:23-43
    std::pair<std::vector<bool>, std::vector<Error>> Assume(const std::vector<std::string>& nodes, const Pod& pod, const PolicySpec& spec,
                                                            bool isLoadSchedule);
    std::vector<int> Score(const std::vector<std::string>& nodes, const Pod& pod, const PolicySpec& spec, bool isLoadSchedule);
    Error Bind(const std::string& node, Pod* pod, const PolicySpec& spec, bool isLoadSchedule);  // pod gets the annotations of pkg/utils/pod.go:65-79
    Error Allocate(const Pod& pod);
    Error Release(const Pod& pod);
    Error Forget(const Pod& pod);
    bool KnownPod(const Pod& pod);
    bool PodReleased(const Pod& pod);
    std::string PrintStatus(const Pod& pod, const std::string& action);  // returns what the reference logs
    std::map<std::string, std::vector<int>> Status();                    // node -> free percent per GPU
    std::pair<std::map<int, GPUCoreUsage>, bool> GetCoreUsage(const std::string& node);
    std::pair<std::map<int, GPUMemoryUsage>, bool> GetMemoryUsage(const std::string& node);
    std::pair<std::map<int, GPUMemoryUsage>, bool> GetMemoryUsageLock(const std::string& node) { return GetMemoryUsage(node); }
    std::pair<std::map<int, GPUCoreUsage>, bool> GetCoreUsageLock(const std::string& node) { return GetCoreUsage(node); }
    void AddCoreUsage(const std::string& node);
    void AddMemoryUsage(const std::string& node);
    void UpdateCoreUsage(const std::string& node, const std::string& coreUsage, const std::string& updateTime, int cardNum);
    void UpdateMemoryUsage(const std::string& node, const std::string& memoryUsage, const std::string& updateTime, int cardNum);
    // GetUsage (nodeusage.go:82-111): (exist, value, error); `now_ns` replaces time.Now()
    struct Usage {
        bool exist;
        double value;
        Error err;
    };
    Usage GetUsage(const std::string& node, const std::string& key, int card, int64_t activeDuration_ns);

    // time.Now() of the next calls (tests freeze it); 0 = wall clock
    void SetNow(int64_t unix_ns) { now_ns_ = unix_ns; }

  private:
    ClusterSource* src_;
    RaterKind rater_;
    ngs_dealer* h_ = nullptr;
    std::string init_error_;
    int64_t now_ns_ = 0;
    std::map<std::string, uint32_t> index_;      // NodeMaps keys -> dense row
    std::vector<std::string> names_;             // row -> name
    std::vector<int> gpu_count_;                 // row -> G
    std::map<std::string, Pod> pod_maps_;        // PodMaps (dealer.go:82)
    std::set<std::string> released_;             // ReleasedPodMap (dealer.go:86)
    std::map<std::string, std::map<int, GPUCoreUsage>> core_usage_;      // dealer.go:84
    std::map<std::string, std::map<int, GPUMemoryUsage>> memory_usage_;  // dealer.go:85

    int policy() const;
    Error getNodeInfo(const std::string& name, uint32_t* row);  // dealer.go:271-301
    bool load_ctx(const PolicySpec& spec, bool isLoad, ngs_load_ctx* out) const;
    int64_t now() const;
};

// helpers mirrored from the reference (used by the tests, too)
std::vector<int32_t> NewDemandFromPod(const Pod& pod);                                   // allocate.go:54-62
Error NewPlanFromPod(const Pod& pod, std::vector<int32_t>* demand, std::vector<int32_t>* gpu_idx);  // allocate.go:29-50
std::string DemandString(const std::vector<int32_t>& d);   // how fmt renders a Demand value with %s: "[(20) (40)]"
std::string GPUsString(const std::vector<int>& free);      // GPUs.String() (allocate.go:133-139): "(10)(50)"
std::string IntsString(const std::vector<int32_t>& v);     // fmt %v of []int: "[1 0]"
int GetGPUDeviceCountOfNode(const Node& node);              // pkg/utils/node.go:8-14
bool IsAssumed(const Pod& pod);                             // pkg/utils/pod.go:81-83
Error getActiveDuration(const std::vector<Period>& list, const std::string& name, int64_t* out_ns);  // stats.go:57-66

}  // namespace ngs_host
