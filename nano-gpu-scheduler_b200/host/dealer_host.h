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
#include <functional>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <string_view>
#include <unordered_map>
#include <memory>
#include <utility>
#include <vector>

#include "../../include/ngs_dealer.h"

namespace ngs_host {

constexpr int kRowCells = NGS_MAX_GPUS_WIDE;  // cells per row in the host-side copies of snapshot rows

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
    bool completed = false;  // utils.IsCompletedPod: DeletionTimestamp set, or phase Succeeded / Failed
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
    virtual std::vector<Pod> ListAssumedPods(const std::string& /*node_name*/) { return {}; }    // label nano-gpu/assume=true
    virtual bool GetPod(const std::string& /*ns*/, const std::string& /*name*/, Pod* /*out*/) { return false; }  // clientset Pods(ns).Get; false: `pods "x" not found`
    virtual std::vector<Pod> ListAllAssumedPods() { return {}; }                              // the same label over all nodes (dealer.go:58)
};

// An error is a string; "" is Go's nil.
using Error = std::string;

// name -> row, open addressing over a name arena; replaces the Go map lookups of dealer.go:272
class NodeIndex {
  public:
    NodeIndex();
    const uint32_t* Find(std::string_view name) const;
    // rows[i] = the row of names[i], 0xFFFFFFFF when absent.  Looks the names up in blocks with the table
    // and arena lines prefetched, so a 100k-name list does not pay one cache miss after the other.
    void FindMany(const std::string_view* names, size_t n, uint32_t* rows) const;
    void Insert(std::string_view name, uint32_t row);
    size_t size() const { return count_; }

  private:
    struct Slot {
        uint64_t hash;
        uint32_t row, len;
        uint32_t off, used;  // used == 0: empty
    };
    std::vector<Slot> slots_;
    std::string arena_;
    size_t count_ = 0, mask_ = 0;
    void grow();
};
uint64_t HashName(std::string_view s);


// A candidate list after name resolution (dealer.go:97-106 done once per list instead of once per
// goroutine): rows[i] = the dense row of names[i], 0xFFFFFFFF when getNodeInfo failed.
struct CandidateSet {
    std::vector<std::string_view> names;     // caller-owned storage
    std::vector<uint32_t> rows;
    std::map<uint32_t, Error> lookup_err;    // i -> getNodeInfo's error, only for the failed ones
    size_t n_missing = 0;
    bool all_rows = false;                   // names are exactly rows 0..n-1 in order: the C ABI's idx == NULL form
};
// an array that is NOT value-initialised: the plans of 100k candidates are 1.2 MB of which only the infeasible nodes' rows are
// ever written or read (ngs_assume_failed_plans), and zero-filling it cost more than the device call
template <class T>
struct RawArray {
    std::unique_ptr<T[]> p;
    size_t n = 0;
    void resize(size_t m) {
        if (m > n) p.reset(new T[m]);
        n = m;
    }
    T* data() { return p.get(); }
    const T* data() const { return p.get(); }
    const T* begin() const { return p.get(); }
    T& operator[](size_t i) { return p[i]; }
    const T& operator[](size_t i) const { return p[i]; }
};
// what one Assume over a CandidateSet brings back from the device
struct AssumeResult {
    std::vector<uint8_t> ok;                 // Assume's []bool
    std::vector<int8_t> reason;              // NGS_REASON_*
    std::vector<int32_t> demand;
    RawArray<int32_t> plan, n_emitted;       // valid for the candidates with reason == NGS_REASON_INFEASIBLE only
    std::string demand_text;                 // DemandString(demand), shared by every failure message
    std::vector<int32_t> free;               // rows [free_first, free_first + free.size()/kRowCells) read back once, for the messages
    uint32_t free_first = 0;
    Error call_err;                          // the C ABI call itself failed: every node reports it
};

enum class RaterKind { Binpack, Spread, Sample };  // controller.Rater (pkg/controller/controller.go:38)

struct GPUCoreUsage {
    std::string CoreUsage, UpdateTime;
};  // nodeusage.go:10-13
struct GPUMemoryUsage {
    std::string MemoryUsage, UpdateTime;
};  // nodeusage.go:15-18


class Dealer {
  public:
    // NewDealer (dealer.go:45-75)
    Dealer(ClusterSource* src, RaterKind rater, uint32_t max_nodes, int device = 0);
    ~Dealer();
    bool ok() const { return h_ != nullptr; }
    ClusterSource* source() const { return src_; }  // the lister / apiserver stand-in the routes fetch pods from
    const std::string& init_error() const { return init_error_; }
    // Assume / Score / Bind as requests to the library's resident kernel instead of one launch each (ngs_serve_verbs): same
    // results, about a third less time per pod; the kernel keeps the GPU's SMs occupied while it waits.
    Error UseResidentKernel(bool on);

    // the 19 methods of dealer.Dealer, same order as dealer.go:23-43
    std::pair<std::vector<bool>, std::vector<Error>> Assume(const std::vector<std::string>& nodes, const Pod& pod, const PolicySpec& spec,
                                                            bool isLoadSchedule);
    std::vector<int> Score(const std::vector<std::string>& nodes, const Pod& pod, const PolicySpec& spec, bool isLoadSchedule);
    // pod gets the annotations of pkg/utils/pod.go:65-79.  `commit` stands where the reference talks to the apiserver
    // (Pods().Update, one retry on the optimistic-lock message, then Pods().Bind: dealer.go:177-199): it is called with
    // the annotated pod after the capacity has been taken; if it returns an error, Bind returns it and - as in the
    // reference - the capacity stays taken while the pod does NOT enter PodMaps (dealer.go:200 is never reached).
    using CommitFn = std::function<Error(const Pod&)>;
    Error Bind(const std::string& node, Pod* pod, const PolicySpec& spec, bool isLoadSchedule, const CommitFn& commit = nullptr);
    Error Allocate(const Pod& pod);
    Error Release(const Pod& pod);
    Error Forget(const Pod& pod);
    bool KnownPod(const Pod& pod);
    bool PodReleased(const Pod& pod);
    std::string PrintStatus(const Pod& pod, const std::string& action);  // returns what the reference logs
    std::map<std::string, std::vector<int>> Status();                    // node -> free percent per GPU
    // what StatusRoute writes (pkg/routes/routes.go:212-240): json.Marshal(map[string]*NodeInfo) with NodeInfo =
    // {Rater, Name, GPUs []{Percent, PercentTotal, RemainLoad}, PlanCache} (node.go:18-23, allocate.go:141-145).
    // PlanCache is written as {}: the per-node plan cache is not kept as objects here (Score / Bind re-derive the
    // plan from the row), so the cached entry the reference would show for the last filtered demand is absent.
    std::string StatusJSON();
    std::pair<std::map<int, GPUCoreUsage>, bool> GetCoreUsage(const std::string& node);
    std::pair<std::map<int, GPUMemoryUsage>, bool> GetMemoryUsage(const std::string& node);
    std::pair<std::map<int, GPUMemoryUsage>, bool> GetMemoryUsageLock(const std::string& node) {  // nodeusage.go:41-46
        std::lock_guard<std::recursive_mutex> lock(mu_);
        return GetMemoryUsage(node);
    }
    std::pair<std::map<int, GPUCoreUsage>, bool> GetCoreUsageLock(const std::string& node) {  // nodeusage.go:34-39
        std::lock_guard<std::recursive_mutex> lock(mu_);
        return GetCoreUsage(node);
    }
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

    // The same Assume / Score over a pre-resolved candidate list (extender_codec.h keeps one per
    // distinct NodeNames array): no per-name strings, one ranged read-back for the failure texts.
    void Resolve(CandidateSet* cs);  // getNodeInfo for every name not resolved yet; new nodes go up in one bulk load
    AssumeResult AssumeSet(const CandidateSet& cs, const Pod& pod, const PolicySpec& spec, bool isLoadSchedule);
    Error FailureText(const CandidateSet& cs, const AssumeResult& r, size_t i) const;  // res[i] of Assume for a node with !ok[i]
    // FailureText for the common case (the rater said no) written at p without temporaries; FailureIsPlain
    // tells whether entry i is that case.  The text is ASCII without quotes or backslashes.
    bool FailureIsPlain(const CandidateSet& cs, const AssumeResult& r, size_t i) const {
        return r.call_err.empty() && r.reason[i] == NGS_REASON_INFEASIBLE && (cs.n_missing == 0 || !cs.lookup_err.count((uint32_t)i));
    }
    char* PutFailureText(char* p, const CandidateSet& cs, const AssumeResult& r, size_t i) const;
    std::vector<int> ScoreSet(const CandidateSet& cs, const Pod& pod, const PolicySpec& spec, bool isLoadSchedule);

    // Controller.annotatorNode (pkg/controller/node.go:111-135): one Prometheus sample for (node, metric key, card)
    // becomes Add*Usage (first sample of the node) + Update*Usage stamped with the controller's local time.
    // `value` is the string QueryLasterData returned (FormatMetricValue); empty -> the reference's error.
    Error AnnotatorNode(const std::string& node, const std::string& key, const std::string& cardNum, const std::string& value);

    // time.Now() of the next calls (tests freeze it); 0 = wall clock
    void SetNow(int64_t unix_ns) { now_ns_ = unix_ns; }

  private:
    // DealerImpl.Lock (dealer.go:81): every method the reference locks takes it here, too, which also gives the
    // "one call at a time per handle" the C library asks for.  Recursive because Assume / Score go through the
    // *Set forms, which lock on their own for callers such as ExtenderRoutes.
    std::recursive_mutex mu_;
    ClusterSource* src_;
    RaterKind rater_;
    ngs_dealer* h_ = nullptr;
    std::string init_error_;
    int64_t now_ns_ = 0;
    NodeIndex index_;                            // NodeMaps keys -> dense row
    std::vector<std::string> names_;             // row -> name
    std::vector<int> gpu_count_;                 // row -> G
    std::map<std::string, Pod> pod_maps_;        // PodMaps (dealer.go:82)
    std::set<std::string> released_;             // ReleasedPodMap (dealer.go:86)
    std::map<std::string, std::map<int, GPUCoreUsage>> core_usage_;      // dealer.go:84
    std::map<std::string, std::map<int, GPUMemoryUsage>> memory_usage_;  // dealer.go:85

    const std::unordered_map<std::string, std::vector<const Pod*>>* listed_ = nullptr;  // NewDealer's one list, grouped by node

    int policy() const;
    Error getNodeInfo(const std::string& name, uint32_t* row);  // dealer.go:271-301
    bool read_rows(uint32_t lo, uint32_t n, std::vector<int32_t>* fr, std::vector<int8_t>* rl) const;  // kRowCells per row
    void replay_assumed_pods(const std::string& name, int G, int32_t* row_free);  // dealer.go:287-299 on the host copy of a new row
    void replay_usage(const std::string& name, uint32_t row);                     // usage samples that arrived before the node existed
    bool load_ctx(const PolicySpec& spec, bool isLoad, ngs_load_ctx* out) const;
    int64_t now() const;
};

// encoding/json's string encoder (defined in extender_codec.cpp, declared again in extender_codec.h)
void AppendJSONString(std::string* out, std::string_view s);

// The controller's pod event handlers, the callers of Allocate / Release / Forget (pkg/controller/controller.go).
// SyncPod: Controller.syncPod (:210-245) for the pod the lister returned (nullptr = NotFound: nothing to do).
struct SyncResult {
    bool forget;  // the work-queue item is done
    Error err;
};
SyncResult SyncPod(Dealer& d, const Pod* pod);
bool NeedUpdate(Dealer& d, const Pod& oldPod, const Pod& newPod);  // updatePodInCache (:289-335): does the update go on the queue?
void DeletePodFromCache(Dealer& d, const Pod& pod);                // deletePodFromCache (:337-357)

// getSyncPeriodFormPolicyConfig (pkg/controller/controller.go:359-371): spec.syncPeriod of the policy file
// (deploy/dynamic-scheduler-node-annotator-cm.yaml:7-16) -> PolicySpec.  The file is YAML; this reads the subset
// the policy uses (block mappings and sequences, scalars, comments) and Go duration strings ("15s", "1m30s",
// "500ms").  Like the reference, a file that does not parse yields an empty list (the error is only logged).
PolicySpec ParsePolicyConfig(const std::string& yaml_text, Error* err = nullptr);
bool ParseGoDuration(const std::string& s, int64_t* out_ns);  // time.ParseDuration

// helpers mirrored from the reference (used by the tests, too)
std::vector<int32_t> NewDemandFromPod(const Pod& pod);                                   // allocate.go:54-62
Error NewPlanFromPod(const Pod& pod, std::vector<int32_t>* demand, std::vector<int32_t>* gpu_idx);  // allocate.go:29-50
std::string DemandString(const std::vector<int32_t>& d);   // how fmt renders a Demand value with %s: "[(20) (40)]"
std::string GPUsString(const std::vector<int>& free);      // GPUs.String() (allocate.go:133-139): "(10)(50)"
std::string IntsString(const std::vector<int32_t>& v);     // fmt %v of []int: "[1 0]"
int GetGPUDeviceCountOfNode(const Node& node);              // pkg/utils/node.go:8-14
std::string FormatLocalTime(int64_t unix_ns);               // Controller.getLocalTime (controller/node.go:137-141): Asia/Shanghai wall clock as "2006-01-02T15:04:05Z"
std::string FormatMetricValue(double v);                    // prometheus.go:57-62: negative or NaN -> 0, then FormatFloat(v, 'f', 5, 64)
bool IsAssumed(const Pod& pod);                             // pkg/utils/pod.go:81-83
uint64_t GetGPUPercentFromPodResource(const Pod& pod);      // pkg/utils/pod.go:50-58: sum of the containers' gpu-percent limits
bool IsGPUSharingPod(const Pod& pod);                       // pkg/utils/pod.go:27-29: the informer's filter for pods this dealer tracks
Error getActiveDuration(const std::vector<Period>& list, const std::string& name, int64_t* out_ns);  // stats.go:57-66

}  // namespace ngs_host
