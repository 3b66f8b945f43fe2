// extender_codec_c.cpp — a flat C surface over extender_codec.h for the Python tests and the route
// bench (ctypes cannot call C++).  Test plumbing, not part of the drop-in boundary: the boundary is
// include/ngs_dealer.h, and a Go maintainer would keep encoding/json on their side of it.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "extender_codec.h"

using namespace ngs_host;

namespace {
char* dup_out(const std::string& s, size_t* n) {
    char* p = (char*)malloc(s.size() + 1);
    memcpy(p, s.data(), s.size());
    p[s.size()] = 0;
    if (n) *n = s.size();
    return p;
}

// a synthetic lister: nodes "node-000000".."node-(n-1)" with the GPU counts given at creation
class RangeSource : public ClusterSource {
  public:
    std::vector<uint8_t> g;
    std::map<std::string, std::vector<Pod>> assumed;
    static bool parse(const std::string& name, uint32_t* i) {
        if (name.size() != 11 || name.compare(0, 5, "node-") != 0) return false;
        uint32_t v = 0;
        for (size_t k = 5; k < 11; k++) {
            if (name[k] < '0' || name[k] > '9') return false;
            v = v * 10 + (uint32_t)(name[k] - '0');
        }
        *i = v;
        return true;
    }
    bool GetNode(const std::string& name, Node* out) override {
        uint32_t i;
        if (!parse(name, &i) || i >= g.size()) return false;
        out->name = name;
        out->has_capacity = true;
        out->gpu_percent_capacity = 100 * (int64_t)g[i];
        return true;
    }
    std::vector<Pod> ListAssumedPods(const std::string& node) override {
        auto it = assumed.find(node);
        return it == assumed.end() ? std::vector<Pod>() : it->second;
    }
    std::map<std::string, Pod> pods;  // pending pods by "ns/name"
    bool GetPod(const std::string& ns, const std::string& name, Pod* out) override {
        auto it = pods.find(ns + "/" + name);
        if (it == pods.end()) return false;
        *out = it->second;
        return true;
    }
    std::vector<Pod> ListAllAssumedPods() override {
        std::vector<Pod> all;
        for (auto& kv : assumed) all.insert(all.end(), kv.second.begin(), kv.second.end());
        return all;
    }
};

struct Routes {
    RangeSource src;
    Dealer* dealer = nullptr;
    ExtenderRoutes* routes = nullptr;
    ~Routes() {
        delete routes;
        delete dealer;
    }
};
}  // namespace

extern "C" {

void ngsx_free(char* p) { free(p); }

// decode a filter / prioritize body and hand back what was understood, as JSON:
// {"err":..,"has_pod":..,"has_nodes":..,"has_node_names":..,"node_names":[..],"pod":{..}}
char* ngsx_decode_args(const char* body, size_t n, size_t* out_n) {
    ExtenderArgs a;
    const Error e = DecodeExtenderArgs(std::string_view(body, n), &a);
    std::string o = "{\"err\":";
    AppendJSONString(&o, e);
    o += a.has_pod ? ",\"has_pod\":true" : ",\"has_pod\":false";
    o += a.has_nodes ? ",\"has_nodes\":true" : ",\"has_nodes\":false";
    o += a.has_node_names ? ",\"has_node_names\":true" : ",\"has_node_names\":false";
    o += ",\"node_names\":[";
    for (size_t i = 0; i < a.node_names.size(); i++) {
        if (i) o += ",";
        AppendJSONString(&o, a.node_names[i]);
    }
    o += "],\"pod\":{\"namespace\":";
    AppendJSONString(&o, a.pod.ns);
    o += ",\"name\":";
    AppendJSONString(&o, a.pod.name);
    o += ",\"uid\":";
    AppendJSONString(&o, a.pod.uid);
    o += ",\"nodeName\":";
    AppendJSONString(&o, a.pod.node_name);
    o += ",\"containers\":[";
    for (size_t i = 0; i < a.pod.containers.size(); i++) {
        const Container& c = a.pod.containers[i];
        if (i) o += ",";
        o += "{\"name\":";
        AppendJSONString(&o, c.name);
        o += c.has_gpu_limit ? ",\"has\":true,\"percent\":" : ",\"has\":false,\"percent\":";
        o += std::to_string(c.gpu_percent) + "}";
    }
    o += "]";
    for (int m = 0; m < 2; m++) {
        o += m == 0 ? ",\"annotations\":{" : ",\"labels\":{";
        bool first = true;
        for (auto& kv : (m == 0 ? a.pod.annotations : a.pod.labels)) {
            if (!first) o += ",";
            first = false;
            AppendJSONString(&o, kv.first);
            o += ":";
            AppendJSONString(&o, kv.second);
        }
        o += "}";
    }
    o += "}}";
    return dup_out(o, out_n);
}

char* ngsx_decode_binding(const char* body, size_t n, size_t* out_n) {
    BindingArgs b;
    const Error e = DecodeBindingArgs(std::string_view(body, n), &b);
    std::string o = "{\"err\":";
    AppendJSONString(&o, e);
    o += ",\"PodName\":";
    AppendJSONString(&o, b.PodName);
    o += ",\"PodNamespace\":";
    AppendJSONString(&o, b.PodNamespace);
    o += ",\"PodUID\":";
    AppendJSONString(&o, b.PodUID);
    o += ",\"Node\":";
    AppendJSONString(&o, b.Node);
    o += "}";
    return dup_out(o, out_n);
}

char* ngsx_json_string(const char* s, size_t n, size_t* out_n) {
    std::string o;
    AppendJSONString(&o, std::string_view(s, n));
    return dup_out(o, out_n);
}

int ngsx_quantity(const char* s, size_t n, int64_t* out) { return ParseQuantityValue(std::string_view(s, n), out) ? 1 : 0; }

// blob = the strings back to back, lens[i] their lengths: n_can names, then n_failed (key, message) pairs
char* ngsx_encode_filter(const char* blob, const uint32_t* lens, uint32_t n_can, uint32_t n_failed, size_t* out_n) {
    std::vector<std::string_view> can;
    std::vector<std::pair<std::string_view, std::string>> failed;
    size_t off = 0, k = 0;
    for (uint32_t i = 0; i < n_can; i++, k++) {
        can.emplace_back(blob + off, lens[k]);
        off += lens[k];
    }
    for (uint32_t i = 0; i < n_failed; i++) {
        std::string_view key(blob + off, lens[k]);
        off += lens[k++];
        std::string msg(blob + off, lens[k]);
        off += lens[k++];
        failed.emplace_back(key, msg);
    }
    return dup_out(EncodeFilterResult(can, failed), out_n);
}
char* ngsx_encode_filter_error(const char* s, size_t n, size_t* out_n) { return dup_out(EncodeFilterError(std::string(s, n)), out_n); }
char* ngsx_encode_binding_result(const char* s, size_t n, size_t* out_n) { return dup_out(EncodeBindingResult(std::string(s, n)), out_n); }
char* ngsx_encode_priorities(const char* blob, const uint32_t* lens, const int32_t* scores, uint32_t n, size_t* out_n) {
    std::vector<std::string_view> hosts;
    std::vector<int> sc(scores, scores + n);
    size_t off = 0;
    for (uint32_t i = 0; i < n; i++) {
        hosts.emplace_back(blob + off, lens[i]);
        off += lens[i];
    }
    return dup_out(EncodeHostPriorityList(hosts, sc), out_n);
}

// the policy file -> "name=period_ns;name=period_ns" (or "!error text")
char* ngsx_parse_policy(const char* text, size_t n, size_t* out_n) {
    Error e;
    const PolicySpec spec = ParsePolicyConfig(std::string(text, n), &e);
    std::string o;
    if (!e.empty()) o = "!" + e;
    for (const Period& p : spec.sync_period) o += (o.empty() ? "" : ";") + p.name + "=" + std::to_string(p.period_ns);
    return dup_out(o, out_n);
}
int ngsx_parse_duration(const char* s, size_t n, int64_t* out) { return ParseGoDuration(std::string(s, n), out) ? 1 : 0; }

// NodeIndex self-check: insert n generated names, look all of them (and n absent ones) up; 0 = ok
int ngsx_node_index_check(uint32_t n) {
    NodeIndex ix;
    char buf[64];
    for (uint32_t i = 0; i < n; i++) {
        const int len = snprintf(buf, sizeof buf, "gpu-node-%u.cluster-%u.example", i, i % 7);
        ix.Insert(std::string_view(buf, (size_t)len), i * 3 + 1);
    }
    if (ix.size() != n) return 1;
    for (uint32_t i = 0; i < n; i++) {
        int len = snprintf(buf, sizeof buf, "gpu-node-%u.cluster-%u.example", i, i % 7);
        const uint32_t* r = ix.Find(std::string_view(buf, (size_t)len));
        if (!r || *r != i * 3 + 1) return 2;
        len = snprintf(buf, sizeof buf, "gpu-node-%u.cluster-%u.examplf", i, i % 7);
        if (ix.Find(std::string_view(buf, (size_t)len))) return 3;
    }
    ix.Insert("gpu-node-0.cluster-0.example", 77);  // overwrite keeps one entry
    if (ix.size() != n || *ix.Find("gpu-node-0.cluster-0.example") != 77) return 4;
    return 0;
}

// ---- the two routes over a device-resident dealer (needs a GPU) --------------------------------
// rater: 0 binpack, 1 spread, 2 sample.  gpu_count[n_nodes].  assumed pods are added before creation
// with ngsx_routes_new + ngsx_routes_add_assumed + ngsx_routes_start.
void* ngsx_routes_new(uint32_t n_nodes, const uint8_t* gpu_count) {
    Routes* r = new Routes();
    r->src.g.assign(gpu_count, gpu_count + n_nodes);
    return r;
}
void ngsx_routes_add_assumed(void* h, uint32_t node, const char* uid, const int32_t* percent, const int32_t* gpu_idx, int n_containers) {
    Routes* r = (Routes*)h;
    char name[16];
    snprintf(name, sizeof name, "node-%06u", node);
    Pod p;
    p.ns = "default";
    p.name = uid;
    p.uid = uid;
    p.node_name = name;
    p.annotations["nano-gpu/assume"] = "true";
    for (int k = 0; k < n_containers; k++) {
        Container c;
        c.name = "c" + std::to_string(k);
        c.has_gpu_limit = true;
        c.gpu_percent = percent[k];
        p.containers.push_back(c);
        p.annotations["nano-gpu/container-" + c.name] = std::to_string(gpu_idx[k]);
    }
    r->src.assumed[name].push_back(p);
}
int ngsx_routes_start(void* h, int rater, uint32_t max_nodes, int device, int load_schedule, char** err) {
    Routes* r = (Routes*)h;
    r->dealer = new Dealer(&r->src, rater == 0 ? RaterKind::Binpack : rater == 1 ? RaterKind::Spread : RaterKind::Sample, max_nodes, device);
    if (!r->dealer->ok()) {
        if (err) *err = dup_out(r->dealer->init_error(), nullptr);
        return 1;
    }
    PolicySpec spec;
    spec.sync_period.push_back(Period{"gpu_core_usage_avg", 15LL * 1000000000LL});
    spec.sync_period.push_back(Period{"gpu_memory_usage_avg", 15LL * 1000000000LL});
    r->routes = new ExtenderRoutes(r->dealer, spec, load_schedule != 0);
    return 0;
}
void ngsx_routes_destroy(void* h) { delete (Routes*)h; }
char* ngsx_routes_predicate(void* h, const char* body, size_t n, size_t* out_n) {
    return dup_out(((Routes*)h)->routes->Predicate(std::string_view(body, n)), out_n);
}
char* ngsx_routes_prioritize(void* h, const char* body, size_t n, size_t* out_n, int* panicked) {
    bool p = false;
    std::string o = ((Routes*)h)->routes->Prioritize(std::string_view(body, n), &p);
    if (panicked) *panicked = p ? 1 : 0;
    return dup_out(o, out_n);
}
void ngsx_routes_counters(void* h, uint64_t out[7]) {
    const ExtenderRoutes::Counters& c = ((Routes*)h)->routes->counters();
    out[0] = c.requests;
    out[1] = c.cache_hits;
    out[2] = c.all_rows;
    out[3] = c.ns_decode;
    out[4] = c.ns_resolve;
    out[5] = c.ns_dealer;
    out[6] = c.ns_encode;
}
// a pending pod the bind route can fetch (what the apiserver would return for Pods(ns).Get(name))
void ngsx_routes_add_pod(void* h, const char* ns, const char* name, const char* uid, const int32_t* percent, int n_containers, int completed) {
    Routes* r = (Routes*)h;
    Pod p;
    p.ns = ns;
    p.name = name;
    p.uid = uid;
    p.completed = completed != 0;
    for (int k = 0; k < n_containers; k++) {
        Container c;
        c.name = "c" + std::to_string(k);
        c.has_gpu_limit = true;
        c.gpu_percent = percent[k];
        p.containers.push_back(c);
    }
    r->src.pods[p.ns + "/" + p.name] = p;
}
char* ngsx_routes_bind_route(void* h, const char* body, size_t n, size_t* out_n, int* http_status) {
    return dup_out(((Routes*)h)->routes->Bind(std::string_view(body, n), http_status), out_n);
}
// StatusRoute's body (routes.go:212-240)
char* ngsx_routes_status(void* h, size_t* out_n) { return dup_out(((Routes*)h)->dealer->StatusJSON(), out_n); }
// Dealer.Bind through the mirror (what the bind route calls after fetching the pod): 0 = nil error
int ngsx_routes_bind(void* h, const char* node, const char* uid, const int32_t* percent, int n_containers, int32_t* gpu_idx_out, char** err) {
    Routes* r = (Routes*)h;
    Pod p;
    p.ns = "default";
    p.name = uid;
    p.uid = uid;
    for (int k = 0; k < n_containers; k++) {
        Container c;
        c.name = "c" + std::to_string(k);
        c.has_gpu_limit = true;
        c.gpu_percent = percent[k];
        p.containers.push_back(c);
    }
    PolicySpec spec;
    const Error e = r->dealer->Bind(node, &p, spec, false);
    if (!e.empty()) {
        if (err) *err = dup_out(e, nullptr);
        return 1;
    }
    for (int k = 0; k < n_containers; k++) gpu_idx_out[k] = atoi(p.annotations["nano-gpu/container-c" + std::to_string(k)].c_str());
    return 0;
}

}  // extern "C"
