// dealer_host.cpp — see dealer_host.h.  Mirrors DealerImpl (pkg/dealer/dealer.go, nodeusage.go) on
// top of the C ABI; paths below are relative to /root/reference/pkg/dealer/.
#include "dealer_host.h"

#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <unordered_map>

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

uint64_t GetGPUPercentFromPodResource(const Pod& pod) {
    uint64_t sum = 0;
    for (const Container& c : pod.containers)
        if (c.has_gpu_limit) sum += (uint64_t)c.gpu_percent;  // uint(val.Value())
    return sum;
}
bool IsGPUSharingPod(const Pod& pod) { return GetGPUPercentFromPodResource(pod) > 0; }

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
std::string FormatLocalTime(int64_t unix_ns) {
    // the zone has been UTC+8 without daylight saving since 1991, and the parser (stats.go:35-38) falls back to a fixed +8 too
    int64_t sec = unix_ns / 1000000000LL;
    if (unix_ns % 1000000000LL < 0) sec--;
    const time_t t = (time_t)(sec + 8 * 3600);
    struct tm tmv;
    gmtime_r(&t, &tmv);
    char buf[32];
    strftime(buf, sizeof buf, "%Y-%m-%dT%H:%M:%SZ", &tmv);
    return buf;
}
std::string FormatMetricValue(double v) {
    if (v < 0 || v != v) v = 0;
    char buf[64];
    snprintf(buf, sizeof buf, "%.5f", v);
    return buf;
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

// NodeIndex
uint64_t HashName(std::string_view s) {
    // 8 bytes at a time, multiply-xorshift mixing (node names are short and share long prefixes)
    const uint64_t k = 0x9E3779B97F4A7C15ull;
    uint64_t h = 0xB2005EEDull ^ (s.size() * k);
    const char* p = s.data();
    size_t n = s.size();
    while (n >= 8) {
        uint64_t w;
        memcpy(&w, p, 8);
        h = (h ^ w) * k;
        h ^= h >> 29;
        p += 8;
        n -= 8;
    }
    if (n) {
        uint64_t w = 0;
        memcpy(&w, p, n);
        h = (h ^ w) * k;
        h ^= h >> 29;
    }
    h *= 0xBF58476D1CE4E5B9ull;
    h ^= h >> 32;
    return h;
}

NodeIndex::NodeIndex() {
    slots_.assign(1024, Slot{0, 0, 0, 0, 0});
    mask_ = 1023;
}

void NodeIndex::grow() {
    std::vector<Slot> old;
    old.swap(slots_);
    const size_t cap = old.size() * 2;
    slots_.assign(cap, Slot{0, 0, 0, 0, 0});
    mask_ = cap - 1;
    for (const Slot& o : old) {
        if (!o.used) continue;
        size_t j = o.hash & mask_;
        while (slots_[j].used) j = (j + 1) & mask_;
        slots_[j] = o;
    }
}

const uint32_t* NodeIndex::Find(std::string_view name) const {
    const uint64_t h = HashName(name);
    for (size_t j = h & mask_;; j = (j + 1) & mask_) {
        const Slot& s = slots_[j];
        if (!s.used) return nullptr;
        if (s.hash == h && s.len == name.size() && memcmp(arena_.data() + s.off, name.data(), name.size()) == 0)
            return s.row == 0xFFFFFFFFu ? nullptr : &s.row;  // 0xFFFFFFFF: a name whose row was withdrawn
    }
}

void NodeIndex::FindMany(const std::string_view* names, size_t n, uint32_t* rows) const {
    constexpr size_t B = 32;
    uint64_t h[B];
    for (size_t base = 0; base < n; base += B) {
        const size_t m = n - base < B ? n - base : B;
        for (size_t k = 0; k < m; k++) {
            h[k] = HashName(names[base + k]);
            __builtin_prefetch(&slots_[h[k] & mask_]);
        }
        for (size_t k = 0; k < m; k++) {  // the home slot almost always decides: fetch its name bytes too
            const Slot& s = slots_[h[k] & mask_];
            if (s.used && s.hash == h[k]) __builtin_prefetch(arena_.data() + s.off);
        }
        for (size_t k = 0; k < m; k++) {
            const std::string_view name = names[base + k];
            uint32_t row = 0xFFFFFFFFu;
            for (size_t j = h[k] & mask_;; j = (j + 1) & mask_) {
                const Slot& s = slots_[j];
                if (!s.used) break;
                if (s.hash == h[k] && s.len == name.size() && memcmp(arena_.data() + s.off, name.data(), name.size()) == 0) {
                    row = s.row;
                    break;
                }
            }
            rows[base + k] = row;
        }
    }
}

void NodeIndex::Insert(std::string_view name, uint32_t row) {
    if ((count_ + 1) * 2 > slots_.size()) grow();
    const uint64_t h = HashName(name);
    size_t j = h & mask_;
    for (; slots_[j].used; j = (j + 1) & mask_) {
        Slot& s = slots_[j];
        if (s.hash == h && s.len == name.size() && memcmp(arena_.data() + s.off, name.data(), name.size()) == 0) {
            s.row = row;
            return;
        }
    }
    slots_[j] = Slot{h, row, (uint32_t)name.size(), (uint32_t)arena_.size(), 1};
    arena_.append(name.data(), name.size());
    count_++;
}

// ---- policy file --------------------------------------------------------------------------------
bool ParseGoDuration(const std::string& str, int64_t* out_ns) {  // time.ParseDuration: [-+]?([0-9]*(\.[0-9]*)?[a-zµ]+)+
    size_t i = 0;
    const size_t n = str.size();
    if (n == 0) return false;
    bool neg = false;
    if (str[i] == '+' || str[i] == '-') neg = str[i++] == '-';
    if (i == n) return false;
    if (str.compare(i, std::string::npos, "0") == 0) {
        *out_ns = 0;
        return true;
    }
    long double total = 0;
    while (i < n) {
        long double v = 0, scale = 1;
        bool digits = false;
        while (i < n && str[i] >= '0' && str[i] <= '9') v = v * 10 + (str[i++] - '0'), digits = true;
        if (i < n && str[i] == '.') {
            i++;
            while (i < n && str[i] >= '0' && str[i] <= '9') scale /= 10, v += (str[i++] - '0') * scale, digits = true;
        }
        if (!digits) return false;
        size_t u = i;
        while (i < n && str[i] != '.' && !(str[i] >= '0' && str[i] <= '9')) i++;
        const std::string unit = str.substr(u, i - u);
        long double mult;
        if (unit == "ns") mult = 1;
        else if (unit == "us" || unit == "\xc2\xb5s" || unit == "\xce\xbcs") mult = 1e3L;  // us, µs (U+00B5), μs (U+03BC)
        else if (unit == "ms") mult = 1e6L;
        else if (unit == "s") mult = 1e9L;
        else if (unit == "m") mult = 60e9L;
        else if (unit == "h") mult = 3600e9L;
        else return false;
        total += v * mult;
    }
    if (total > 9.2e18L) return false;
    *out_ns = (int64_t)(neg ? -total : total);
    return true;
}

PolicySpec ParsePolicyConfig(const std::string& text, Error* err) {
    PolicySpec spec;
    if (err) err->clear();
    // lines -> (indent, content without comment); the only structure needed: spec: / syncPeriod: / "- name:" / "period:"
    struct Line {
        int indent;
        std::string body;
    };
    std::vector<Line> lines;
    size_t pos = 0;
    while (pos <= text.size()) {
        size_t e = text.find('\n', pos);
        if (e == std::string::npos) e = text.size();
        std::string ln = text.substr(pos, e - pos);
        pos = e + 1;
        bool in_s = false, in_d = false;
        for (size_t k = 0; k < ln.size(); k++) {  // a comment starts at an unquoted '#' that begins the line or follows a blank
            if (ln[k] == '\'' && !in_d) in_s = !in_s;
            else if (ln[k] == '"' && !in_s) in_d = !in_d;
            else if (ln[k] == '#' && !in_s && !in_d && (k == 0 || ln[k - 1] == ' ' || ln[k - 1] == '\t')) {
                ln.erase(k);
                break;
            }
        }
        while (!ln.empty() && (ln.back() == ' ' || ln.back() == '\t' || ln.back() == '\r')) ln.pop_back();
        size_t ind = 0;
        while (ind < ln.size() && ln[ind] == ' ') ind++;
        if (ind < ln.size() && ln[ind] == '\t') {
            if (err) *err = "yaml: found character that cannot start any token";
            return PolicySpec();
        }
        if (ind == ln.size()) continue;
        lines.push_back({(int)ind, ln.substr(ind)});
    }
    auto scalar = [](std::string v) {
        while (!v.empty() && v.front() == ' ') v.erase(0, 1);
        if (v.size() >= 2 && ((v.front() == '"' && v.back() == '"') || (v.front() == '\'' && v.back() == '\''))) v = v.substr(1, v.size() - 2);
        return v;
    };
    int spec_ind = -1, sp_ind = -1, item_ind = -1;
    bool in_spec = false, in_sp = false;
    for (const Line& L : lines) {
        if (in_sp && L.indent <= sp_ind && L.body.compare(0, 1, "-") != 0) in_sp = false;  // left the sequence
        if (in_sp && L.indent < sp_ind) in_sp = false;
        if (in_spec && L.indent <= spec_ind) in_spec = in_sp = false;
        if (!in_spec) {
            if (L.indent == 0 && (L.body == "spec:" || L.body.compare(0, 5, "spec:") == 0)) {
                in_spec = true;
                spec_ind = L.indent;
            }
            continue;
        }
        if (!in_sp) {
            if (L.body == "syncPeriod:") {
                in_sp = true;
                sp_ind = L.indent;
                item_ind = -1;
            }
            continue;
        }
        std::string body = L.body;
        int ind = L.indent;
        if (body[0] == '-') {  // a new sequence entry: "- name: x"
            spec.sync_period.emplace_back();
            size_t k = 1;
            while (k < body.size() && body[k] == ' ') k++;
            ind += (int)k;
            body = body.substr(k);
            item_ind = ind;
            if (body.empty()) continue;
        } else if (spec.sync_period.empty() || ind != item_ind) {
            continue;  // deeper or foreign structure: not part of a Period
        }
        const size_t c = body.find(':');
        if (c == std::string::npos) continue;
        const std::string key = body.substr(0, c), val = scalar(body.substr(c + 1));
        if (key == "name") spec.sync_period.back().name = val;
        else if (key == "period") {
            int64_t ns = 0;
            if (!ParseGoDuration(val, &ns)) {  // yaml.v2: cannot unmarshal into time.Duration -> the whole Unmarshal reports an error
                if (err) *err = "yaml: unmarshal errors:\n  cannot unmarshal !!str `" + val + "` into time.Duration";
                return PolicySpec();
            }
            spec.sync_period.back().period_ns = ns;
        }
    }
    return spec;
}

// ---- the controller's pod event handlers -------------------------------------------------------
SyncResult SyncPod(Dealer& d, const Pod* pod) {  // controller.go:210-245
    if (!pod) return {true, ""};                 // "has been deleted": logged only
    if (pod->completed) {                        // utils.IsCompletedPod
        d.Release(*pod);                         // an error is logged, not returned (:228-230)
        d.PrintStatus(*pod, "release");
        return {true, ""};
    }
    if (pod->node_name.empty()) return {true, ""};
    const Error e = d.Allocate(*pod);
    d.PrintStatus(*pod, "allocate");
    if (!e.empty()) return {false, e};
    return {true, ""};
}
bool NeedUpdate(Dealer& d, const Pod& oldPod, const Pod& newPod) {  // controller.go:303-310
    if (d.KnownPod(oldPod) && newPod.completed) return true;       // 1. turned complete or failed
    if (!d.KnownPod(oldPod) && !d.PodReleased(oldPod) && IsAssumed(newPod)) return true;  // 2. unknown, unreleased, GPU annotation set
    return false;
}
void DeletePodFromCache(Dealer& d, const Pod& pod) { d.Forget(pod); }  // controller.go:356

// ---- Dealer -----------------------------------------------------------------------------------
Dealer::Dealer(ClusterSource* src, RaterKind rater, uint32_t max_nodes, int device) : src_(src), rater_(rater) {
    int rc = ngs_create(&h_, device, max_nodes, 0, nullptr);
    if (rc != NGS_OK) {
        init_error_ = h_ ? ngs_last_error(h_) : ngs_strerror(rc);
        if (init_error_.empty()) init_error_ = ngs_strerror(rc);
        if (h_) ngs_destroy(h_);
        h_ = nullptr;
        return;
    }
    if (src_) {  // dealer.go:58-72: every node that already runs an assumed pod is built up front
        // One list call serves every node here (the reference lists again per node, dealer.go:279): the
        // pods are grouped by node and replay_assumed_pods reads the groups instead of asking the source.
        const std::vector<Pod> pods = src_->ListAllAssumedPods();
        std::unordered_map<std::string, std::vector<const Pod*>> by_node;
        std::vector<const std::string*> nodes;
        for (const Pod& p : pods) {
            if (p.node_name.empty()) continue;
            auto ins = by_node.try_emplace(p.node_name);
            if (ins.second) nodes.push_back(&ins.first->first);
            ins.first->second.push_back(&p);
        }
        if (!nodes.empty()) {
            CandidateSet cs;
            for (const std::string* s : nodes) cs.names.push_back(*s);
            listed_ = &by_node;
            Resolve(&cs);  // failures are only logged by the reference (dealer.go:68)
            listed_ = nullptr;
        }
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
        if (out->n_periods >= NGS_MAX_PERIODS) {
            // allocate.go:175-195 sums every entry; the kernels hold NGS_MAX_PERIODS of them.  Never truncate silently:
            // an over-long list travels on as n_periods > NGS_MAX_PERIODS, which every C-ABI call refuses (NGS_ERR_UNSUPPORTED)
            out->n_periods = NGS_MAX_PERIODS + 1;
            break;
        }
        out->metric[out->n_periods] = p.name == kCoreUsagePriority ? NGS_METRIC_CORE : NGS_METRIC_MEMORY;  // nodeusage.go:84
        out->window_ns[out->n_periods] = dur;
        out->n_periods++;
    }
    out->now_ns = now();
    return true;
}

Error Dealer::UseResidentKernel(bool on) {
    std::lock_guard<std::recursive_mutex> lock(mu_);
    if (!h_) return init_error_;
    return ngs_serve_verbs(h_, on ? 1 : 0) == NGS_OK ? "" : Error(ngs_last_error(h_));
}

Error Dealer::getNodeInfo(const std::string& name, uint32_t* row) {  // dealer.go:271-301
    if (const uint32_t* known = index_.Find(name)) {
        *row = *known;
        return "";
    }
    CandidateSet one;  // same path as a whole candidate list, so a node is built one way only
    one.names.push_back(name);
    Resolve(&one);
    if (one.n_missing) return one.lookup_err.begin()->second;
    *row = one.rows[0];
    return "";
}

// Rows [lo, lo + n) with all their cells, kRowCells per row (nodes of 9..12 GPUs keep cells 8..11 in a second table:
// ngs_node_get_wide).  rl may be nullptr.
bool Dealer::read_rows(uint32_t lo, uint32_t n, std::vector<int32_t>* fr, std::vector<int8_t>* rl) const {
    fr->assign((size_t)n * kRowCells, 0);
    if (rl) rl->assign((size_t)n * kRowCells, 0);
    if (n == 0) return true;
    std::vector<int32_t> f8((size_t)n * NGS_MAX_GPUS);
    std::vector<int8_t> r8(rl ? (size_t)n * NGS_MAX_GPUS : 0);
    if (ngs_snapshot_read(h_, lo, n, nullptr, f8.data(), rl ? r8.data() : nullptr) != NGS_OK) return false;
    for (uint32_t i = 0; i < n; i++) {
        if (gpu_count_[lo + i] > NGS_MAX_GPUS) {
            if (ngs_node_get_wide(h_, lo + i, nullptr, fr->data() + (size_t)i * kRowCells, rl ? rl->data() + (size_t)i * kRowCells : nullptr) != NGS_OK)
                return false;
            continue;
        }
        for (int g = 0; g < NGS_MAX_GPUS; g++) {
            (*fr)[(size_t)i * kRowCells + g] = f8[(size_t)i * NGS_MAX_GPUS + g];
            if (rl) (*rl)[(size_t)i * kRowCells + g] = r8[(size_t)i * NGS_MAX_GPUS + g];
        }
    }
    return true;
}

// dealer.go:287-299 on the host copy of a new row: Allocate each assumed pod's plan (all or nothing,
// DESIGN.md "deviations"), remember the pods that fitted.
void Dealer::replay_assumed_pods(const std::string& name, int G, int32_t* fr) {
    std::vector<Pod> fetched;
    std::vector<const Pod*> local;
    const std::vector<const Pod*>* pods = &local;
    if (listed_) {
        auto it = listed_->find(name);
        if (it != listed_->end()) pods = &it->second;
    } else {
        fetched = src_->ListAssumedPods(name);
        for (const Pod& p : fetched) local.push_back(&p);
    }
    for (const Pod* pp : *pods) {
        const Pod& pod = *pp;
        std::vector<int32_t> demand, idx;
        if (!NewPlanFromPod(pod, &demand, &idx).empty()) continue;
        if (demand.size() > NGS_MAX_CONTAINERS) continue;
        int32_t f[kRowCells];
        for (int g = 0; g < kRowCells; g++) f[g] = fr[g];
        bool ok = true;
        for (size_t k = 0; k < demand.size() && ok; k++) {
            const int gi = idx[k];
            if (gi < 0) continue;  // allocate.go:105-107
            if (demand[k] < 0 || gi >= G || f[gi] < demand[k]) ok = false;  // allocate.go:108
            else f[gi] -= demand[k];
        }
        if (!ok) continue;
        for (int g = 0; g < kRowCells; g++) fr[g] = f[g];
        pod_maps_[pod.uid] = pod;
    }
}

// usage samples that arrived before the node was materialised
void Dealer::replay_usage(const std::string& name, uint32_t r) {
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
}

void Dealer::Resolve(CandidateSet* cs) {  // dealer.go:97-106 for a whole list
    std::lock_guard<std::recursive_mutex> lock(mu_);
    const size_t n = cs->names.size();
    const bool fresh = cs->rows.size() != n;
    if (fresh) {
        cs->rows.resize(n);
        cs->lookup_err.clear();
        index_.FindMany(cs->names.data(), n, cs->rows.data());  // the known names, which is nearly all of them
    }
    // pass 1: known names; unknown ones are fetched from the lister and queued for one bulk upload
    const uint32_t first_new = (uint32_t)names_.size();
    std::vector<uint8_t> new_g;
    std::vector<int32_t> new_free;
    for (size_t i = 0; i < n; i++) {
        if (cs->rows[i] != 0xFFFFFFFFu) continue;
        const std::string_view nm = cs->names[i];
        if (const uint32_t* known = index_.Find(nm)) {
            cs->rows[i] = *known;
            cs->lookup_err.erase((uint32_t)i);
            continue;
        }
        const std::string name(nm);
        Node node;
        Error e;
        int G = 0;
        if (!src_ || !src_->GetNode(name, &node)) e = fmt("node \"%s\" not found", name.c_str());
        else if ((G = GetGPUDeviceCountOfNode(node)) <= 0) e = fmt("node \"%s\" has no nano-gpu/gpu-percent capacity", name.c_str());
        else if (G > NGS_MAX_GPUS_WIDE) e = fmt("node \"%s\": %d GPUs, more than NGS_MAX_GPUS_WIDE", name.c_str(), G);
        if (!e.empty()) {
            cs->lookup_err[(uint32_t)i] = e;
            continue;
        }
        const uint32_t r = (uint32_t)names_.size();
        index_.Insert(nm, r);
        names_.push_back(name);
        gpu_count_.push_back(G);
        new_g.push_back((uint8_t)G);
        new_free.resize(new_free.size() + kRowCells, 0);
        int32_t* fr = new_free.data() + new_free.size() - kRowCells;
        for (int g = 0; g < G; g++) fr[g] = 100;  // NewNodeInfo (node.go:25-42)
        replay_assumed_pods(name, G, fr);
        cs->rows[i] = r;
        cs->lookup_err.erase((uint32_t)i);
    }
    if (!new_g.empty()) {
        // all new rows go up in one H2D copy instead of one launch per node and per replayed pod; the bulk arrays are
        // NGS_MAX_GPUS cells wide, so a node with more GPUs than that travels as an empty row and is set on its own after
        const size_t nn = new_g.size();
        std::vector<uint8_t> g8(nn);
        std::vector<int32_t> f8(nn * NGS_MAX_GPUS);
        for (size_t k = 0; k < nn; k++) {
            const bool wide = new_g[k] > NGS_MAX_GPUS;
            g8[k] = wide ? 0 : new_g[k];
            for (int g = 0; g < NGS_MAX_GPUS; g++) f8[k * NGS_MAX_GPUS + g] = wide ? 0 : new_free[k * kRowCells + g];
        }
        int rc = ngs_snapshot_load(h_, first_new, (uint32_t)nn, g8.data(), f8.data(), nullptr);
        for (size_t k = 0; k < nn && rc == NGS_OK; k++)
            if (new_g[k] > NGS_MAX_GPUS) rc = ngs_node_set_wide(h_, first_new + (uint32_t)k, new_g[k], new_free.data() + k * kRowCells, nullptr);
        if (rc != NGS_OK) {
            const Error e = ngs_last_error(h_);
            for (size_t i = 0; i < n; i++)
                if (cs->rows[i] != 0xFFFFFFFFu && cs->rows[i] >= first_new) cs->rows[i] = 0xFFFFFFFFu, cs->lookup_err[(uint32_t)i] = e;
            // the rows were never built: forget them again
            for (uint32_t r = first_new; r < names_.size(); r++) index_.Insert(names_[r], 0xFFFFFFFFu);
            names_.resize(first_new);
            gpu_count_.resize(first_new);
        } else {
            for (uint32_t r = first_new; r < names_.size(); r++) replay_usage(names_[r], r);
        }
    }
    cs->n_missing = cs->lookup_err.size();
    cs->all_rows = cs->n_missing == 0 && n == names_.size() && n > 0;
    for (size_t i = 0; cs->all_rows && i < n; i++) cs->all_rows = cs->rows[i] == i;
}

AssumeResult Dealer::AssumeSet(const CandidateSet& cs, const Pod& pod, const PolicySpec& spec, bool isLoadSchedule) {  // dealer.go:89-136
    std::lock_guard<std::recursive_mutex> lock(mu_);
    AssumeResult r;
    r.demand = NewDemandFromPod(pod);
    r.demand_text = DemandString(r.demand);
    const size_t n = cs.names.size();
    r.ok.assign(n, 0);
    r.reason.assign(n, NGS_REASON_UNKNOWN_NODE);
    if (n == 0) return r;
    ngs_load_ctx L;
    const bool has_load = load_ctx(spec, isLoadSchedule, &L);
    const int C = (int)r.demand.size();
    r.plan.resize(n * (size_t)(C > 0 ? C : 1));
    r.n_emitted.resize(n);
    // (plans only where a FailedNodes text will quote them: the infeasible candidates' partial `indexes`)
    const int rc = ngs_assume_failed_plans(h_, cs.all_rows ? nullptr : cs.rows.data(), (uint32_t)n, r.demand.data(), C, policy(), has_load ? &L : nullptr,
                                           r.ok.data(), r.reason.data(), nullptr, r.plan.data(), r.n_emitted.data());
    if (rc != NGS_OK) {
        r.call_err = ngs_last_error(h_);
        std::fill(r.ok.begin(), r.ok.end(), 0);
        return r;
    }
    // one ranged read-back covers every row a failure message will quote
    uint32_t lo = 0xFFFFFFFFu, hi = 0;
    for (size_t i = 0; i < n; i++)
        if (r.reason[i] == NGS_REASON_INFEASIBLE) {
            lo = std::min(lo, cs.rows[i]);
            hi = std::max(hi, cs.rows[i]);
        }
    if (lo <= hi) {
        r.free_first = lo;
        read_rows(lo, hi - lo + 1, &r.free, nullptr);
    }
    return r;
}

Error Dealer::FailureText(const CandidateSet& cs, const AssumeResult& r, size_t i) const {
    auto it = cs.lookup_err.find((uint32_t)i);
    if (it != cs.lookup_err.end()) return "nano gpu scheduler get node failed: " + it->second;  // dealer.go:102
    if (!r.call_err.empty()) return r.call_err;
    if (r.reason[i] != NGS_REASON_INFEASIBLE) return "";
    const uint32_t row = cs.rows[i];
    const int32_t* fr = r.free.data() + (size_t)(row - r.free_first) * kRowCells;
    std::vector<int> gp(fr, fr + gpu_count_[row]);
    if (rater_ == RaterKind::Sample)  // rater.go:47
        return fmt("unexpected failure of allocation  %s on %s", DemandString(r.demand).c_str(), GPUsString(gp).c_str());
    const size_t C = r.demand.size();  // rater.go:99,152
    std::vector<int32_t> partial(r.plan.begin() + i * C, r.plan.begin() + i * C + r.n_emitted[i]);
    return fmt("can't allocate   %s on %s: indexes=%s", DemandString(r.demand).c_str(), GPUsString(gp).c_str(), IntsString(partial).c_str());
}

char* Dealer::PutFailureText(char* p, const CandidateSet& cs, const AssumeResult& r, size_t i) const {
    auto lit = [&](const char* s) {
        const size_t n = strlen(s);
        memcpy(p, s, n);
        p += n;
    };
    auto num = [&](int64_t v) {
        char tmp[24];
        int n = 0;
        uint64_t u = v < 0 ? (uint64_t)0 - (uint64_t)v : (uint64_t)v;
        do tmp[n++] = (char)('0' + u % 10); while (u /= 10);
        if (v < 0) *p++ = '-';
        while (n) *p++ = tmp[--n];
    };
    const uint32_t row = cs.rows[i];
    const int32_t* fr = r.free.data() + (size_t)(row - r.free_first) * kRowCells;
    lit(rater_ == RaterKind::Sample ? "unexpected failure of allocation  " : "can't allocate   ");  // rater.go:47 / 99,152
    memcpy(p, r.demand_text.data(), r.demand_text.size());
    p += r.demand_text.size();
    lit(" on ");
    for (int g = 0; g < gpu_count_[row]; g++) {  // GPUs.String() (allocate.go:133-139)
        *p++ = '(';
        num(fr[g]);
        *p++ = ')';
    }
    if (rater_ != RaterKind::Sample) {
        lit(": indexes=[");
        const size_t C = r.demand.size();
        for (int k = 0; k < r.n_emitted[i]; k++) {
            if (k) *p++ = ' ';
            num(r.plan[i * C + (size_t)k]);
        }
        *p++ = ']';
    }
    return p;
}

std::vector<int> Dealer::ScoreSet(const CandidateSet& cs, const Pod& pod, const PolicySpec& spec, bool isLoadSchedule) {  // dealer.go:138-153
    std::lock_guard<std::recursive_mutex> lock(mu_);
    const std::vector<int32_t> demand = NewDemandFromPod(pod);
    const size_t n = cs.names.size();
    std::vector<int> scores(n, NGS_SCORE_MIN);  // a missing node scores ScoreMin (dealer.go:145-148)
    if (n == 0) return scores;
    ngs_load_ctx L;
    const bool has_load = load_ctx(spec, isLoadSchedule, &L);
    const int32_t* view = nullptr;  // the scores where the kernels left them: one copy (into the result), not two
    if (ngs_score_view(h_, cs.all_rows ? nullptr : cs.rows.data(), (uint32_t)n, demand.data(), (int)demand.size(), policy(), has_load ? &L : nullptr,
                       &view) == NGS_OK)
        std::copy(view, view + n, scores.begin());
    return scores;
}

std::pair<std::vector<bool>, std::vector<Error>> Dealer::Assume(const std::vector<std::string>& nodes, const Pod& pod,
                                                                const PolicySpec& spec, bool isLoadSchedule) {  // dealer.go:89-136
    std::lock_guard<std::recursive_mutex> lock(mu_);
    CandidateSet cs;
    cs.names.assign(nodes.begin(), nodes.end());
    Resolve(&cs);
    const AssumeResult r = AssumeSet(cs, pod, spec, isLoadSchedule);
    const size_t n = nodes.size();
    std::vector<bool> ans(n, false);
    std::vector<Error> res(n);
    for (size_t i = 0; i < n; i++) {
        ans[i] = r.ok[i] != 0;
        if (!ans[i]) res[i] = FailureText(cs, r, i);
    }
    return {ans, res};
}

std::vector<int> Dealer::Score(const std::vector<std::string>& nodes, const Pod& pod, const PolicySpec& spec, bool isLoadSchedule) {
    std::lock_guard<std::recursive_mutex> lock(mu_);
    CandidateSet cs;
    cs.names.assign(nodes.begin(), nodes.end());
    Resolve(&cs);
    return ScoreSet(cs, pod, spec, isLoadSchedule);
}

Error Dealer::Bind(const std::string& node, Pod* pod, const PolicySpec& spec, bool isLoadSchedule, const CommitFn& commit) {  // dealer.go:155-203
    std::lock_guard<std::recursive_mutex> lock(mu_);
    uint32_t row;
    Error e = getNodeInfo(node, &row);
    if (!e.empty()) return e;
    const std::vector<int32_t> demand = NewDemandFromPod(*pod);
    ngs_load_ctx L;
    const bool has_load = load_ctx(spec, isLoadSchedule, &L);
    std::vector<int32_t> gpu(demand.size() ? demand.size() : 1);
    int rc = ngs_bind(h_, row, demand.data(), (int)demand.size(), policy(), has_load ? &L : nullptr, gpu.data(), nullptr);
    if (rc == NGS_ERR_INFEASIBLE) {  // node.go:75
        std::vector<int32_t> fr;
        read_rows(row, 1, &fr, nullptr);
        std::vector<int> gp(fr.begin(), fr.begin() + gpu_count_[row]);
        return fmt("assume %s on %s failed", DemandString(demand).c_str(), GPUsString(gp).c_str());
    }
    if (rc != NGS_OK) return ngs_last_error(h_);
    // GetUpdatedPodAnnotationSpec (pkg/utils/pod.go:65-79); the apiserver Update + Bind of dealer.go:177-199 stay with the host
    for (size_t i = 0; i < pod->containers.size(); i++) pod->annotations[std::string(kContainerOn) + pod->containers[i].name] = std::to_string(gpu[i]);
    pod->annotations[kGPUAssume] = "true";
    pod->labels[kGPUAssume] = "true";
    if (commit) {  // dealer.go:177-199: a failed apiserver Update / Bind returns before PodMaps is written; nothing is rolled back
        Error ce = commit(*pod);
        if (!ce.empty()) return ce;
    }
    pod_maps_[pod->uid] = *pod;  // dealer.go:200
    return "";
}

Error Dealer::Allocate(const Pod& pod) {  // dealer.go:205-228
    std::lock_guard<std::recursive_mutex> lock(mu_);
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
        std::vector<int32_t> fr;
        read_rows(row, 1, &fr, nullptr);
        std::vector<int> gp(fr.begin(), fr.begin() + gpu_count_[row]);
        return fmt("can't apply plan &{%s %s 0} on %s", DemandString(demand).c_str(), IntsString(idx).c_str(), GPUsString(gp).c_str());
    }
    pod_maps_[pod.uid] = pod;
    return "";
}

Error Dealer::Release(const Pod& pod) {  // dealer.go:230-255
    std::lock_guard<std::recursive_mutex> lock(mu_);
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
    std::lock_guard<std::recursive_mutex> lock(mu_);
    released_.erase(pod.uid);
    pod_maps_.erase(pod.uid);
    return "";
}
bool Dealer::KnownPod(const Pod& pod) {  // dealer.go:257-262
    std::lock_guard<std::recursive_mutex> lock(mu_);
    return pod_maps_.count(pod.uid) != 0;
}
bool Dealer::PodReleased(const Pod& pod) {  // dealer.go:264-269
    std::lock_guard<std::recursive_mutex> lock(mu_);
    return released_.count(pod.uid) != 0;
}

std::map<std::string, std::vector<int>> Dealer::Status() {  // dealer.go:321-323 (NodeMaps), via one read-back
    std::lock_guard<std::recursive_mutex> lock(mu_);
    std::map<std::string, std::vector<int>> out;
    const uint32_t n = (uint32_t)names_.size();
    if (n == 0) return out;
    std::vector<int32_t> fr;
    if (!read_rows(0, n, &fr, nullptr)) return out;
    for (uint32_t r = 0; r < n; r++) out[names_[r]] = std::vector<int>(fr.begin() + (size_t)r * kRowCells, fr.begin() + (size_t)r * kRowCells + gpu_count_[r]);
    return out;
}

std::string Dealer::StatusJSON() {  // routes.go:212-240
    std::lock_guard<std::recursive_mutex> lock(mu_);
    const uint32_t n = (uint32_t)names_.size();
    std::vector<int32_t> fr;
    std::vector<int8_t> rl;
    if (!read_rows(0, n, &fr, &rl)) return "{}";
    std::vector<uint32_t> order(n);
    for (uint32_t r = 0; r < n; r++) order[r] = r;
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return names_[a] < names_[b]; });  // Go sorts map keys
    std::string out = "{";
    for (uint32_t k = 0; k < n; k++) {
        const uint32_t r = order[k];
        if (k) out += ",";
        AppendJSONString(&out, names_[r]);
        out += ":{\"Rater\":{},\"Name\":";
        AppendJSONString(&out, names_[r]);
        out += ",\"GPUs\":[";
        for (int g = 0; g < gpu_count_[r]; g++) {
            if (g) out += ",";
            out += "{\"Percent\":" + std::to_string(fr[(size_t)r * kRowCells + g]) + ",\"PercentTotal\":100,\"RemainLoad\":" +
                   std::to_string((int)rl[(size_t)r * kRowCells + g]) + "}";
        }
        out += "],\"PlanCache\":{}}";
    }
    return out + "}";
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
    std::lock_guard<std::recursive_mutex> lock(mu_);
    core_usage_[node] = {};
    if (const uint32_t* row = index_.Find(node)) ngs_load_add_node(h_, *row, NGS_METRIC_CORE);
}
void Dealer::AddMemoryUsage(const std::string& node) {  // nodeusage.go:64-68
    std::lock_guard<std::recursive_mutex> lock(mu_);
    memory_usage_[node] = {};
    if (const uint32_t* row = index_.Find(node)) ngs_load_add_node(h_, *row, NGS_METRIC_MEMORY);
}
void Dealer::UpdateCoreUsage(const std::string& node, const std::string& coreUsage, const std::string& updateTime, int cardNum) {
    std::lock_guard<std::recursive_mutex> lock(mu_);
    core_usage_[node][cardNum] = GPUCoreUsage{coreUsage, updateTime};  // nodeusage.go:70-74
    const uint32_t* row = index_.Find(node);
    if (!row) return;  // replayed when the node is materialised (getNodeInfo)
    double v = 0;
    int64_t ts = 0;
    int vok = ngs_parse_usage(coreUsage.c_str(), &v), tok = ngs_parse_update_time(updateTime.c_str(), &ts);
    ngs_load_update(h_, *row, NGS_METRIC_CORE, cardNum, v, vok, ts, tok);
}
void Dealer::UpdateMemoryUsage(const std::string& node, const std::string& memoryUsage, const std::string& updateTime, int cardNum) {
    std::lock_guard<std::recursive_mutex> lock(mu_);
    memory_usage_[node][cardNum] = GPUMemoryUsage{memoryUsage, updateTime};  // nodeusage.go:76-80
    const uint32_t* row = index_.Find(node);
    if (!row) return;
    double v = 0;
    int64_t ts = 0;
    int vok = ngs_parse_usage(memoryUsage.c_str(), &v), tok = ngs_parse_update_time(updateTime.c_str(), &ts);
    ngs_load_update(h_, *row, NGS_METRIC_MEMORY, cardNum, v, vok, ts, tok);
}

Error Dealer::AnnotatorNode(const std::string& node, const std::string& key, const std::string& cardNum, const std::string& value) {
    if (value.empty())  // controller/node.go:113-116
        return fmt("QueryLasterData %s for node %s return value: %s , error: <nil>", key.c_str(), node.c_str(), value.c_str());
    char* end = nullptr;
    const long card = strtol(cardNum.c_str(), &end, 10);  // strconv.Atoi
    if (cardNum.empty() || *end != '\0') return fmt("strconv.Atoi: parsing \"%s\": invalid syntax", cardNum.c_str());
    const std::string stamp = FormatLocalTime(now());
    if (key == kCoreUsagePriority) {
        if (!GetCoreUsageLock(node).second) AddCoreUsage(node);
        UpdateCoreUsage(node, value, stamp, (int)card);
    } else {
        if (!GetMemoryUsageLock(node).second) AddMemoryUsage(node);
        UpdateMemoryUsage(node, value, stamp, (int)card);
    }
    return "";
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
