// dealer_host.cpp — see dealer_host.h.  Mirrors DealerImpl (pkg/dealer/dealer.go, nodeusage.go) on
// top of the C ABI; paths below are relative to /root/reference/pkg/dealer/.
#include "dealer_host.h"

#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

namespace ngs_host {

namespace {
const char* kGPUAssume = "nano-gpu/assume";                    // pkg/types/types.go:12
const char* kContainerOn = "nano-gpu/container-";              // pkg/types/types.go:15 ("nano-gpu/container-%s")
const char* kCoreUsagePriority = "gpu_core_usage_avg";         // type.go:7
constexpr int64_t kExtenderActivePeriodNs = 5LL * 60 * 1000000000LL;  // type.go:6

std::string fmt(const char* f, ...) __attribute__((format(printf, 1, 2)));
std::string fmt(const char* f, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, f);
    vsnprintf(buf, sizeof(buf), f, ap);
    va_end(ap);
    return buf;
}
}  // namespace

// ---- helpers mirrored from the reference ----------------------------------------------------
std::vector<int32_t> NewDemandFromPod(const Pod& pod) {  // allocate.go:54-62
    std::vector<int32_t> d;
    for (const Container& c : pod.containers) d.push_back(c.has_gpu_limit ? (int32_t)c.gpu_percent : 0);
    return d;
}

bool IsAssumed(const Pod& pod) {
    auto it = pod.annotations.find(kGPUAssume);
    return it != pod.annotations.end() && it->second == "true";
}

Error NewPlanFromPod(const Pod& pod, std::vector<int32_t>* demand, std::vector<int32_t>* gpu_idx) {  // allocate.go:29-50
    if (!IsAssumed(pod)) return fmt("pod %s/%s is not assumed", pod.ns.c_str(), pod.name.c_str());
    *demand = NewDemandFromPod(pod);
    gpu_idx->clear();
    for (const Container& c : pod.containers) {
        int idx = 0;  // a missing or unparsable annotation yields index 0 (allocate.go:42-46)
        auto it = pod.annotations.find(std::string(kContainerOn) + c.name);
        if (it != pod.annotations.end()) {
            char* end = nullptr;
            long v = strtol(it->second.c_str(), &end, 10);
            if (end != it->second.c_str() && *end == '\0') idx = (int)v;
        }
        gpu_idx->push_back(idx);
    }
    return "";
}

std::string DemandString(const std::vector<int32_t>& d) {
    // fmt "%s" of a Demand VALUE: String() is declared on *Demand (allocate.go:64), so the slice is
    // printed element-wise with GPUResource.String() (allocate.go:147-149): "[(20) (40)]"
    std::string s = "[";
    for (size_t i = 0; i < d.size(); i++) s += (i ? " (" : "(") + std::to_string(d[i]) + ")";
    return s + "]";
}
std::string GPUsString(const std::vector<int>& free) {  // GPUs.String(), allocate.go:133-139
    std::string s;
    for (int f : free) s += "(" + std::to_string(f) + ")";
    return s;
}
std::string IntsString(const std::vector<int32_t>& v) {  // fmt %v of []int
    std::string s = "[";
    for (size_t i = 0; i < v.size(); i++) s += (i ? " " : "") + std::to_string(v[i]);
    return s + "]";
}
int GetGPUDeviceCountOfNode(const Node& node) {  // pkg/utils/node.go:8-14
    if (!node.has_capacity) return 0;
    return (int)(node.gpu_percent_capacity / 100);
}
Error getActiveDuration(const std::vector<Period>& list, const std::string& name, int64_t* out_ns) {  // stats.go:57-66
    for (const Period& p : list)
        if (p.name == name && p.period_ns != 0) {
            *out_ns = p.period_ns + kExtenderActivePeriodNs;
            return "";
        }
    *out_ns = 0;
    return "error get activeDuration";
}

// ---- Dealer -----------------------------------------------------------------------------------
Dealer::Dealer(ClusterSource* src, RaterKind rater, uint32_t max_nodes, int device) : src_(src), rater_(rater) {
    int rc = ngs_create(&h_, device, max_nodes, 0, nullptr);
    if (rc != NGS_OK) {
        init_error_ = h_ ? ngs_last_error(h_) : ngs_strerror(rc);
        if (init_error_.empty()) init_error_ = ngs_strerror(rc);
        if (h_) ngs_destroy(h_);
        h_ = nullptr;
    }
}
Dealer::~Dealer() {
    if (h_) ngs_destroy(h_);
}

int Dealer::policy() const {
    return rater_ == RaterKind::Binpack ? NGS_POLICY_BINPACK : rater_ == RaterKind::Spread ? NGS_POLICY_SPREAD : NGS_POLICY_RANDOM;
}
int64_t Dealer::now() const {
    if (now_ns_) return now_ns_;
    return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::system_clock::now().time_since_epoch()).count();
}

bool Dealer::load_ctx(const PolicySpec& spec, bool isLoad, ngs_load_ctx* out) const {
    if (!isLoad) return false;
    *out = ngs_load_ctx{};
    for (const Period& p : spec.sync_period) {  // allocate.go:175-180
        int64_t dur = 0;
        if (!getActiveDuration(spec.sync_period, p.name, &dur).empty() || dur == 0) continue;
        if (out->n_periods >= NGS_MAX_PERIODS) break;
        out->metric[out->n_periods] = p.name == kCoreUsagePriority ? NGS_METRIC_CORE : NGS_METRIC_MEMORY;  // nodeusage.go:84
        out->window_ns[out->n_periods] = dur;
        out->n_periods++;
    }
    out->now_ns = now();
    return true;
}

Error Dealer::getNodeInfo(const std::string& name, uint32_t* row) {  // dealer.go:271-301
    auto it = index_.find(name);
    if (it != index_.end()) {
        *row = it->second;
        return "";
    }
    Node node;
    if (!src_ || !src_->GetNode(name, &node)) return fmt("node \"%s\" not found", name.c_str());
    const int G = GetGPUDeviceCountOfNode(node);
    if (G <= 0) return fmt("node \"%s\" has no nano-gpu/gpu-percent capacity", name.c_str());  // the reference would divide by zero in Rate
    const uint32_t r = (uint32_t)names_.size();
    int rc = ngs_node_upsert(h_, r, G);  // NewNodeInfo (node.go:25-42)
    if (rc != NGS_OK) return ngs_last_error(h_);
    index_[name] = r;
    names_.push_back(name);
    gpu_count_.push_back(G);
    for (const Pod& pod : src_->ListAssumedPods(name)) {  // dealer.go:287-299: replay the assumed pods
        std::vector<int32_t> demand, idx;
        if (!NewPlanFromPod(pod, &demand, &idx).empty()) continue;
        if (ngs_node_apply(h_, r, idx.data(), demand.data(), (int)demand.size(), +1) != NGS_OK) continue;
        pod_maps_[pod.uid] = pod;
    }
    // usage samples that arrived before the node was materialised
    for (int m = 0; m < 2; m++) {
        if (m == 0 ? !core_usage_.count(name) : !memory_usage_.count(name)) continue;
        ngs_load_add_node(h_, r, m);
        if (m == 0) {
            for (auto& kv : core_usage_[name]) {
                double v = 0;
                int64_t ts = 0;
                int vok = ngs_parse_usage(kv.second.CoreUsage.c_str(), &v), tok = ngs_parse_update_time(kv.second.UpdateTime.c_str(), &ts);
                ngs_load_update(h_, r, m, kv.first, v, vok, ts, tok);
            }
        } else {
            for (auto& kv : memory_usage_[name]) {
                double v = 0;
                int64_t ts = 0;
                int vok = ngs_parse_usage(kv.second.MemoryUsage.c_str(), &v), tok = ngs_parse_update_time(kv.second.UpdateTime.c_str(), &ts);
                ngs_load_update(h_, r, m, kv.first, v, vok, ts, tok);
            }
        }
    }
    *row = r;
    return "";
}

std::pair<std::vector<bool>, std::vector<Error>> Dealer::Assume(const std::vector<std::string>& nodes, const Pod& pod,
                                                                const PolicySpec& spec, bool isLoadSchedule) {  // dealer.go:89-136
    const std::vector<int32_t> demand = NewDemandFromPod(pod);
    const size_t n = nodes.size();
    std::vector<bool> ans(n, false);
    std::vector<Error> res(n);
    std::vector<uint32_t> idx(n, 0xFFFFFFFFu);
    for (size_t i = 0; i < n; i++) {
        uint32_t row;
        Error e = getNodeInfo(nodes[i], &row);
        if (!e.empty()) res[i] = "nano gpu scheduler get node failed: " + e;  // dealer.go:102
        else idx[i] = row;
    }
    if (n == 0) return {ans, res};
    ngs_load_ctx L;
    const bool has_load = load_ctx(spec, isLoadSchedule, &L);
    const int C = (int)demand.size();
    std::vector<uint8_t> feas(n);
    std::vector<int8_t> reason(n);
    std::vector<int32_t> plan(n * (size_t)(C > 0 ? C : 1)), nem(n);
    int rc = ngs_assume(h_, idx.data(), (uint32_t)n, demand.data(), C, policy(), has_load ? &L : nullptr, feas.data(), reason.data(),
                        nullptr, plan.data(), nem.data());
    for (size_t i = 0; i < n; i++) {
        if (!res[i].empty()) continue;  // node lookup already failed
        if (rc != NGS_OK) {
            res[i] = ngs_last_error(h_);
            continue;
        }
        ans[i] = feas[i] != 0;
        if (reason[i] == NGS_REASON_INFEASIBLE) {
            std::vector<int32_t> fr(8);
            ngs_snapshot_read(h_, idx[i], 1, nullptr, fr.data(), nullptr);
            std::vector<int> gp(fr.begin(), fr.begin() + gpu_count_[idx[i]]);
            if (rater_ == RaterKind::Sample) {  // rater.go:47
                res[i] = fmt("unexpected failure of allocation  %s on %s", DemandString(demand).c_str(), GPUsString(gp).c_str());
            } else {  // rater.go:99,152
                std::vector<int32_t> partial(plan.begin() + i * C, plan.begin() + i * C + nem[i]);
                res[i] = fmt("can't allocate   %s on %s: indexes=%s", DemandString(demand).c_str(), GPUsString(gp).c_str(),
                             IntsString(partial).c_str());
            }
        }
    }
    return {ans, res};
}

std::vector<int> Dealer::Score(const std::vector<std::string>& nodes, const Pod& pod, const PolicySpec& spec, bool isLoadSchedule) {
    const std::vector<int32_t> demand = NewDemandFromPod(pod);  // dealer.go:138-153
    const size_t n = nodes.size();
    std::vector<int> scores(n, NGS_SCORE_MIN);
    if (n == 0) return scores;
    std::vector<uint32_t> idx(n, 0xFFFFFFFFu);
    for (size_t i = 0; i < n; i++) {
        uint32_t row;
        if (getNodeInfo(nodes[i], &row).empty()) idx[i] = row;  // a missing node scores ScoreMin (dealer.go:145-148)
    }
    ngs_load_ctx L;
    const bool has_load = load_ctx(spec, isLoadSchedule, &L);
    std::vector<int32_t> sc(n);
    if (ngs_score(h_, idx.data(), (uint32_t)n, demand.data(), (int)demand.size(), policy(), has_load ? &L : nullptr, sc.data()) == NGS_OK)
        for (size_t i = 0; i < n; i++) scores[i] = sc[i];
    return scores;
}

Error Dealer::Bind(const std::string& node, Pod* pod, const PolicySpec& spec, bool isLoadSchedule) {  // dealer.go:155-203
    uint32_t row;
    Error e = getNodeInfo(node, &row);
    if (!e.empty()) return e;
    const std::vector<int32_t> demand = NewDemandFromPod(*pod);
    ngs_load_ctx L;
    const bool has_load = load_ctx(spec, isLoadSchedule, &L);
    std::vector<int32_t> gpu(demand.size() ? demand.size() : 1);
    int rc = ngs_bind(h_, row, demand.data(), (int)demand.size(), policy(), has_load ? &L : nullptr, gpu.data(), nullptr);
    if (rc == NGS_ERR_INFEASIBLE) {  // node.go:75
        std::vector<int32_t> fr(8);
        ngs_snapshot_read(h_, row, 1, nullptr, fr.data(), nullptr);
        std::vector<int> gp(fr.begin(), fr.begin() + gpu_count_[row]);
        return fmt("assume %s on %s failed", DemandString(demand).c_str(), GPUsString(gp).c_str());
    }
    if (rc != NGS_OK) return ngs_last_error(h_);
    // GetUpdatedPodAnnotationSpec (pkg/utils/pod.go:65-79); the apiserver Update + Bind of dealer.go:177-199 stay with the host
    for (size_t i = 0; i < pod->containers.size(); i++) pod->annotations[std::string(kContainerOn) + pod->containers[i].name] = std::to_string(gpu[i]);
    pod->annotations[kGPUAssume] = "true";
    pod->labels[kGPUAssume] = "true";
    pod_maps_[pod->uid] = *pod;  // dealer.go:200
    return "";
}

Error Dealer::Allocate(const Pod& pod) {  // dealer.go:205-228
    if (pod.node_name.empty()) return fmt("pod %s/%s nodename is empty", pod.ns.c_str(), pod.name.c_str());
    uint32_t row;
    Error e = getNodeInfo(pod.node_name, &row);
    if (!e.empty()) return e;
    if (pod_maps_.count(pod.uid)) return "";
    std::vector<int32_t> demand, idx;
    e = NewPlanFromPod(pod, &demand, &idx);
    if (!e.empty()) return e;
    int rc = ngs_node_apply(h_, row, idx.data(), demand.data(), (int)demand.size(), +1);
    if (rc != NGS_OK) {  // allocate.go:113: "can't apply plan %v on %s" with plan = &Plan{Demand, GPUIndexes, Score}
        std::vector<int32_t> fr(8);
        ngs_snapshot_read(h_, row, 1, nullptr, fr.data(), nullptr);
        std::vector<int> gp(fr.begin(), fr.begin() + gpu_count_[row]);
        return fmt("can't apply plan &{%s %s 0} on %s", DemandString(demand).c_str(), IntsString(idx).c_str(), GPUsString(gp).c_str());
    }
    pod_maps_[pod.uid] = pod;
    return "";
}

Error Dealer::Release(const Pod& pod) {  // dealer.go:230-255
    uint32_t row;
    Error e = getNodeInfo(pod.node_name, &row);
    if (!e.empty()) return e;
    if (!pod_maps_.count(pod.uid)) return "";  // "no such pod": logged, nil
    std::vector<int32_t> demand, idx;
    e = NewPlanFromPod(pod, &demand, &idx);
    if (!e.empty()) return e;
    int rc = ngs_node_apply(h_, row, idx.data(), demand.data(), (int)demand.size(), -1);
    if (rc != NGS_OK) {  // allocate.go:126
        for (int32_t i : idx)
            if (i >= gpu_count_[row]) return fmt("allocate plan's GPU index %d bigger then GPU resource", i);
        return ngs_last_error(h_);
    }
    pod_maps_.erase(pod.uid);
    released_.insert(pod.uid);
    return "";
}

Error Dealer::Forget(const Pod& pod) {  // dealer.go:311-319: bookkeeping only, no capacity returns
    released_.erase(pod.uid);
    pod_maps_.erase(pod.uid);
    return "";
}
bool Dealer::KnownPod(const Pod& pod) { return pod_maps_.count(pod.uid) != 0; }     // dealer.go:257-262
bool Dealer::PodReleased(const Pod& pod) { return released_.count(pod.uid) != 0; }  // dealer.go:264-269

std::map<std::string, std::vector<int>> Dealer::Status() {  // dealer.go:321-323 (NodeMaps), via one read-back
    std::map<std::string, std::vector<int>> out;
    const uint32_t n = (uint32_t)names_.size();
    if (n == 0) return out;
    std::vector<int32_t> fr((size_t)n * 8);
    if (ngs_snapshot_read(h_, 0, n, nullptr, fr.data(), nullptr) != NGS_OK) return out;
    for (uint32_t r = 0; r < n; r++) out[names_[r]] = std::vector<int>(fr.begin() + r * 8, fr.begin() + r * 8 + gpu_count_[r]);
    return out;
}

std::string Dealer::PrintStatus(const Pod& pod, const std::string& action) {  // dealer.go:303-309
    std::string s = fmt("------resource status after %s for %s/%s------\n", action.c_str(), pod.ns.c_str(), pod.name.c_str());
    for (auto& kv : Status()) s += fmt("node %s: %s\n", kv.first.c_str(), GPUsString(kv.second).c_str());
    return s + "------------\n";
}

std::pair<std::map<int, GPUCoreUsage>, bool> Dealer::GetCoreUsage(const std::string& node) {  // nodeusage.go:48-51
    auto it = core_usage_.find(node);
    if (it == core_usage_.end()) return {{}, false};
    return {it->second, true};
}
std::pair<std::map<int, GPUMemoryUsage>, bool> Dealer::GetMemoryUsage(const std::string& node) {  // nodeusage.go:53-56
    auto it = memory_usage_.find(node);
    if (it == memory_usage_.end()) return {{}, false};
    return {it->second, true};
}
void Dealer::AddCoreUsage(const std::string& node) {  // nodeusage.go:58-62
    core_usage_[node] = {};
    auto it = index_.find(node);
    if (it != index_.end()) ngs_load_add_node(h_, it->second, NGS_METRIC_CORE);
}
void Dealer::AddMemoryUsage(const std::string& node) {  // nodeusage.go:64-68
    memory_usage_[node] = {};
    auto it = index_.find(node);
    if (it != index_.end()) ngs_load_add_node(h_, it->second, NGS_METRIC_MEMORY);
}
void Dealer::UpdateCoreUsage(const std::string& node, const std::string& coreUsage, const std::string& updateTime, int cardNum) {
    core_usage_[node][cardNum] = GPUCoreUsage{coreUsage, updateTime};  // nodeusage.go:70-74
    auto it = index_.find(node);
    if (it == index_.end()) return;  // replayed when the node is materialised (getNodeInfo)
    double v = 0;
    int64_t ts = 0;
    int vok = ngs_parse_usage(coreUsage.c_str(), &v), tok = ngs_parse_update_time(updateTime.c_str(), &ts);
    ngs_load_update(h_, it->second, NGS_METRIC_CORE, cardNum, v, vok, ts, tok);
}
void Dealer::UpdateMemoryUsage(const std::string& node, const std::string& memoryUsage, const std::string& updateTime, int cardNum) {
    memory_usage_[node][cardNum] = GPUMemoryUsage{memoryUsage, updateTime};  // nodeusage.go:76-80
    auto it = index_.find(node);
    if (it == index_.end()) return;
    double v = 0;
    int64_t ts = 0;
    int vok = ngs_parse_usage(memoryUsage.c_str(), &v), tok = ngs_parse_update_time(updateTime.c_str(), &ts);
    ngs_load_update(h_, it->second, NGS_METRIC_MEMORY, cardNum, v, vok, ts, tok);
}

Dealer::Usage Dealer::GetUsage(const std::string& node, const std::string& key, int card, int64_t activeDuration_ns) {  // nodeusage.go:82-111
    std::string usage, tm;
    if (key == kCoreUsagePriority) {
        auto it = core_usage_.find(node);
        if (it == core_usage_.end()) return {false, 0, ""};
        auto c = it->second.find(card);
        if (c != it->second.end()) {
            usage = c->second.CoreUsage;
            tm = c->second.UpdateTime;
        }
    } else {
        auto it = memory_usage_.find(node);
        if (it == memory_usage_.end()) return {false, 0, ""};
        auto c = it->second.find(card);
        if (c != it->second.end()) {
            usage = c->second.MemoryUsage;
            tm = c->second.UpdateTime;
        }
    }
    int64_t ts = 0;
    // inUpdateTimePeriod (stats.go:30-55)
    if (!ngs_parse_update_time(tm.c_str(), &ts) || !(now() < ts * 1000000000LL + activeDuration_ns))
        return {true, 0, key + "not in update period"};
    double v = 0;
    if (!ngs_parse_usage(usage.c_str(), &v)) return {true, 0, key + "strconv.ParseFloat error"};
    if (v < 0 || v > 1) return {true, 0, key + " usage < 0 || usage > 1"};
    return {true, v, ""};
}

}  // namespace ngs_host
