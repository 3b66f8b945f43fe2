// test_dealer_host.cpp — the reference's own unit tests, re-run through the C++ mirror of the
// Dealer interface (nano-gpu-scheduler_b200/host) and therefore through the C ABI and the CUDA
// kernels.  Tables are the ones of /root/reference/pkg/dealer/rater_test.go and allocate_test.go
// (cited per test); the Dealer-level sequence tests have no counterpart in the reference, which
// never tests DealerImpl.  Needs a GPU.  Exit code 0 = all passed.
#include <cstdio>
#include <cstdlib>
#include <map>
#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "../../nano-gpu-scheduler_b200/host/dealer_host.h"

using namespace ngs_host;

static int g_fail = 0, g_checks = 0;
#define CHECK(cond)                                                                     \
    do {                                                                                \
        g_checks++;                                                                     \
        if (!(cond)) {                                                                  \
            g_fail++;                                                                   \
            printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);                     \
        }                                                                               \
    } while (0)
#define CHECK_EQ_STR(a, b)                                                              \
    do {                                                                                \
        g_checks++;                                                                     \
        const std::string a_ = (a), b_ = (b);                                           \
        if (a_ != b_) {                                                                 \
            g_fail++;                                                                   \
            printf("FAIL %s:%d:\n  got  \"%s\"\n  want \"%s\"\n", __FILE__, __LINE__, a_.c_str(), b_.c_str()); \
        }                                                                               \
    } while (0)

// informer / apiserver stand-in
struct FakeCluster : ClusterSource {
    std::map<std::string, Node> nodes;
    std::map<std::string, std::vector<Pod>> assumed;
    bool GetNode(const std::string& name, Node* out) override {
        auto it = nodes.find(name);
        if (it == nodes.end()) return false;
        *out = it->second;
        return true;
    }
    std::vector<Pod> ListAssumedPods(const std::string& n) override { return assumed[n]; }
    void add(const std::string& name, int gpus) {
        Node n;
        n.name = name;
        n.has_capacity = true;
        n.gpu_percent_capacity = 100 * gpus;
        nodes[name] = n;
    }
};

// MockPodWithDemand (allocate_test.go:108-122)
static Pod MockPodWithDemand(const std::vector<int>& demand, const std::string& uid = "uid") {
    Pod p;
    p.ns = "default";
    p.name = "pod-" + uid;
    p.uid = uid;
    for (size_t i = 0; i < demand.size(); i++) {
        Container c;
        c.name = std::to_string(i);
        c.has_gpu_limit = true;
        c.gpu_percent = demand[i];
        p.containers.push_back(c);
    }
    return p;
}
// MockPodWithPlan (allocate_test.go:88-106)
static Pod MockPodWithPlan(const std::vector<int>& demand, const std::vector<int>& idx, const std::string& node, const std::string& uid) {
    Pod p = MockPodWithDemand(demand, uid);
    p.node_name = node;
    p.annotations["nano-gpu/assume"] = "true";
    for (size_t i = 0; i < idx.size(); i++) p.annotations["nano-gpu/container-" + std::to_string(i)] = std::to_string(idx[i]);
    return p;
}

// a one-node cluster whose GPUs hold `free` percent: fresh node + one pre-assumed pod taking the rest
static void setup_node(FakeCluster& fc, const std::string& name, const std::vector<int>& free) {
    fc.add(name, (int)free.size());
    std::vector<int> used, idx;
    for (size_t g = 0; g < free.size(); g++)
        if (free[g] != 100) {
            used.push_back(100 - free[g]);
            idx.push_back((int)g);
        }
    if (!used.empty()) fc.assumed[name].push_back(MockPodWithPlan(used, idx, name, "pre-" + name));
}

struct ChooseCase {
    std::vector<int> gpus, demand, expected;
    bool expectFailed;
};

static std::vector<int> plan_of(const Pod& p) {
    std::vector<int> out;
    for (const Container& c : p.containers) out.push_back(atoi(p.annotations.at("nano-gpu/container-" + c.name).c_str()));
    return out;
}

static void run_choose_cases(RaterKind kind, const std::vector<ChooseCase>& cases, const char* what) {
    int k = 0;
    for (const ChooseCase& tc : cases) {
        FakeCluster fc;
        std::string name = std::string(what) + std::to_string(k++);
        setup_node(fc, name, tc.gpus);
        Dealer d(&fc, kind, 16);
        CHECK(d.ok());
        Pod pod = MockPodWithDemand(tc.demand);
        auto r = d.Assume({name}, pod, PolicySpec{}, false);
        CHECK(r.first[0] == !tc.expectFailed);
        CHECK(r.second[0].empty() == !tc.expectFailed);  // ans[i]=false is always paired with an error (predicate.go:30)
        Error e = d.Bind(name, &pod, PolicySpec{}, false);
        CHECK(e.empty() == !tc.expectFailed);
        if (!tc.expectFailed) CHECK(plan_of(pod) == tc.expected);
    }
}

static void TestBinpackChoose() {  // rater_test.go:133-266
    run_choose_cases(RaterKind::Binpack,
                     {{{100, 100}, {20, 40}, {0, 0}, false},
                      {{20, 100}, {20, 40}, {0, 1}, false},
                      {{10, 100}, {20, 40}, {1, 1}, false},
                      {{100, 100}, {0, 40, 40}, {-1, 0, 0}, false},
                      {{10, 50}, {20, 40}, {}, true}},
                     "bp");
}
static void TestSpreadChoose() {  // rater_test.go:268-401
    run_choose_cases(RaterKind::Spread,
                     {{{100, 100}, {20, 40}, {0, 1}, false},
                      {{20, 100}, {20, 40}, {1, 1}, false},
                      {{10, 100}, {20, 40}, {1, 1}, false},
                      {{100, 100}, {0, 40, 40}, {-1, 0, 1}, false},
                      {{10, 50}, {20, 40}, {}, true}},
                     "sp");
}
static void TestChoose() {  // allocate_test.go:160-190 (SampleRater)
    struct C {
        std::vector<int> gpus, demand;
        bool success;
    } cases[] = {{{100, 100}, {50, 50}, true}, {{100}, {50, 50}, true}, {{100}, {50, 60}, false}, {{100, 100}, {100, 10}, true}};
    int k = 0;
    for (auto& tc : cases) {
        FakeCluster fc;
        std::string name = "sm" + std::to_string(k++);
        setup_node(fc, name, tc.gpus);
        Dealer d(&fc, RaterKind::Sample, 16);
        auto r = d.Assume({name}, MockPodWithDemand(tc.demand), PolicySpec{}, false);
        CHECK(r.first[0] == tc.success);
    }
}

static int rate(RaterKind kind, const std::vector<int>& gpus) {
    FakeCluster fc;
    setup_node(fc, "n", gpus);
    Dealer d(&fc, kind, 16);
    return d.Score({"n"}, MockPodWithDemand({0}), PolicySpec{}, false)[0];
}
static void TestBinpackRate() {  // rater_test.go:9-37
    CHECK(rate(RaterKind::Binpack, {30, 50}) < rate(RaterKind::Binpack, {20, 40}));
    CHECK(rate(RaterKind::Binpack, {30, 50}) == 58);
}
static void TestSpreadRate() {  // rater_test.go:39-131
    struct C {
        std::vector<int> g1, g2;
        bool firstIsPreferred;
    } cases[] = {{{30, 50}, {20, 40}, true}, {{90}, {50, 40}, true}, {{100}, {50, 50}, true}, {{100}, {100, 50}, false}};
    for (auto& tc : cases) CHECK((rate(RaterKind::Spread, tc.g1) > rate(RaterKind::Spread, tc.g2)) == tc.firstIsPreferred);
}

static void TestNewDemandAndPlanFromPod() {  // allocate_test.go:124-158
    Pod p = MockPodWithPlan({20, 40, 0}, {1, 0, -1}, "n", "u");
    std::vector<int32_t> d, idx;
    CHECK(NewPlanFromPod(p, &d, &idx).empty());
    CHECK((d == std::vector<int32_t>{20, 40, 0}) && (idx == std::vector<int32_t>{1, 0, -1}));
    CHECK((NewDemandFromPod(MockPodWithDemand({5, 10})) == std::vector<int32_t>{5, 10}));
    CHECK(GetGPUPercentFromPodResource(MockPodWithDemand({5, 10})) == 15 && IsGPUSharingPod(MockPodWithDemand({5, 10})));
    CHECK(!IsGPUSharingPod(MockPodWithDemand({0, 0})) && !IsGPUSharingPod(MockPodWithDemand({})));  // pkg/utils/pod.go:27-29
    Pod q = MockPodWithDemand({20});
    CHECK_EQ_STR(NewPlanFromPod(q, &d, &idx), "pod default/pod-uid is not assumed");
}

// ---- Dealer-level behaviour (the reference has no tests for DealerImpl) ----
static void TestAssumeErrorTexts() {
    FakeCluster fc;
    setup_node(fc, "full", {10, 50});
    Dealer d(&fc, RaterKind::Binpack, 16);
    auto r = d.Assume({"full", "ghost"}, MockPodWithDemand({20, 40}), PolicySpec{}, false);
    CHECK(!r.first[0] && !r.first[1]);
    CHECK_EQ_STR(r.second[0], "can't allocate   [(20) (40)] on (10)(50): indexes=[1]");           // rater.go:99
    CHECK_EQ_STR(r.second[1], "nano gpu scheduler get node failed: node \"ghost\" not found");       // dealer.go:102
    std::vector<int> sc = d.Score({"full", "ghost"}, MockPodWithDemand({20, 40}), PolicySpec{}, false);
    CHECK(sc[0] == 0 && sc[1] == 0);                                                                 // ScoreMin (node.go:64, dealer.go:146)
    Pod p = MockPodWithDemand({20, 40});
    CHECK_EQ_STR(d.Bind("full", &p, PolicySpec{}, false), "assume [(20) (40)] on (10)(50) failed");  // node.go:75
}

// dealer.go:177-200: when the apiserver Update / Bind fails the capacity stays taken and the pod stays out of PodMaps
static void TestBindCommitFailureLeavesPodUnknown() {
    FakeCluster fc;
    setup_node(fc, "n", {100, 100});
    Dealer d(&fc, RaterKind::Binpack, 16);
    Pod p = MockPodWithDemand({20, 40}, "commit-fails");
    p.node_name = "n";
    int calls = 0;
    Error e = d.Bind("n", &p, PolicySpec{}, false, [&](const Pod& annotated) {
        calls++;
        CHECK(annotated.annotations.count("nano-gpu/assume") == 1);  // the host sees the annotated copy it has to write
        return Error("Operation cannot be fulfilled on pods");
    });
    CHECK_EQ_STR(e, "Operation cannot be fulfilled on pods");
    CHECK(calls == 1);
    CHECK(!d.KnownPod(p));                    // never reached dealer.go:200
    auto st = d.Status();
    CHECK(st["n"][0] == 40 && st["n"][1] == 100);  // ... but the capacity is gone (no rollback in the reference)
    CHECK(d.Release(p).empty());              // "no such pod": Release of an unknown pod changes nothing (dealer.go:240-243)
    st = d.Status();
    CHECK(st["n"][0] == 40);
    // five usable SyncPeriod entries are within the library's limit of eight; nine are refused loudly, not truncated
    PolicySpec many;
    for (int i = 0; i < 9; i++) many.sync_period.push_back(Period{i % 2 ? "gpu_core_usage_avg" : "gpu_memory_usage_avg", 15LL * 1000000000LL});
    auto r = d.Assume({"n"}, MockPodWithDemand({10}), many, true);
    CHECK(!r.first[0]);
    CHECK(r.second[0].find("periods") != std::string::npos);
}

static void ScheduleCycleAndRelease(bool resident) {
    FakeCluster fc;
    for (int i = 0; i < 6; i++) fc.add("node-" + std::to_string(i), 8);
    fc.add("small", 2);
    Dealer d(&fc, RaterKind::Binpack, 64);
    if (resident) CHECK(d.UseResidentKernel(true).empty());  // the same cycle, the verbs served by the resident kernel
    std::vector<std::string> names = {"node-0", "node-1", "node-2", "node-3", "node-4", "node-5", "small"};
    std::vector<Pod> bound;
    for (int k = 0; k < 40; k++) {
        Pod pod = MockPodWithDemand({30, 50}, "p" + std::to_string(k));
        auto r = d.Assume(names, pod, PolicySpec{}, false);
        std::vector<int> sc = d.Score(names, pod, PolicySpec{}, false);
        int best = -1;
        for (size_t i = 0; i < names.size(); i++)
            if (r.first[i] && (best < 0 || sc[i] > sc[best])) best = (int)i;
        if (best < 0) break;
        pod.node_name = names[best];
        CHECK(d.Bind(names[best], &pod, PolicySpec{}, false).empty());
        CHECK(d.KnownPod(pod));
        bound.push_back(pod);
    }
    CHECK(bound.size() >= 30);
    // binpack fills a node before it opens the next one: at most one node is partly used
    auto st = d.Status();
    int partly = 0;
    for (auto& kv : st) {
        int touched = 0, untouched = 0;
        for (int f : kv.second) (f == 100 ? untouched : touched)++;
        if (touched && untouched) partly++;  // a node with both used and still-empty GPUs
    }
    CHECK(partly <= 1);
    // Release gives the capacity back, Forget does not (dealer.go:311-319)
    for (Pod& p : bound) {
        CHECK(d.Release(p).empty());
        CHECK(d.PodReleased(p) && !d.KnownPod(p));
    }
    for (auto& kv : d.Status())
        for (int f : kv.second) CHECK(f == 100);
    CHECK(d.Release(bound[0]).empty());  // unknown pod: logged, nil — and nothing is added twice
    for (auto& kv : d.Status())
        for (int f : kv.second) CHECK(f == 100);
    CHECK(d.Forget(bound[0]).empty() && !d.PodReleased(bound[0]));
}

static void TestAllocateFromAnnotationsAndRebuild() {
    FakeCluster fc;
    fc.add("n", 4);
    fc.assumed["n"].push_back(MockPodWithPlan({60, 30}, {0, 0}, "n", "old"));  // found by getNodeInfo's pod list (dealer.go:279-299)
    Dealer d(&fc, RaterKind::Spread, 16);
    CHECK((d.Status().empty()));
    Pod running = MockPodWithPlan({50}, {2}, "n", "new");
    CHECK(d.Allocate(running).empty());
    CHECK((d.Status()["n"] == std::vector<int>{10, 100, 50, 100}));
    CHECK(d.Allocate(running).empty());  // already known: no double accounting (dealer.go:215)
    CHECK((d.Status()["n"] == std::vector<int>{10, 100, 50, 100}));
    Pod too_big = MockPodWithPlan({20}, {0}, "n", "big");
    CHECK_EQ_STR(d.Allocate(too_big), "can't apply plan &{[(20)] [0] 0} on (10)(100)(50)(100)");  // allocate.go:113
    Pod nowhere = MockPodWithDemand({10}, "x");
    CHECK_EQ_STR(d.Allocate(nowhere), "pod default/pod-x nodename is empty");  // dealer.go:209
    CHECK_EQ_STR(d.PrintStatus(running, "allocate"),
                 "------resource status after allocate for default/pod-new------\nnode n: (10)(100)(50)(100)\n------------\n");
}

// a node with 12 GPUs (capacity 1200: pkg/utils/node.go:8-14); expected values from the oracle
static void TestTwelveGpuNode() {
    FakeCluster fc;
    fc.add("wide", 12);
    fc.add("small", 4);
    fc.assumed["wide"].push_back(MockPodWithPlan({60, 100}, {10, 3}, "wide", "old"));  // replayed onto cards 10 and 3 (dealer.go:287-299)
    Dealer d(&fc, RaterKind::Binpack, 16);
    auto r = d.Assume({"wide", "small"}, MockPodWithDemand({50, 50, 30}), PolicySpec{}, false);
    CHECK(r.first[0] && r.first[1]);
    const std::vector<int> start{100, 100, 100, 0, 100, 100, 100, 100, 100, 100, 40, 100};
    CHECK((d.Status()["wide"] == start));
    std::vector<int> sc = d.Score({"wide", "small"}, MockPodWithDemand({50, 50, 30}), PolicySpec{}, false);
    CHECK(sc[0] == 1 && sc[1] == -4);  // int(160/1200*100) - 12; 0 - 4
    // eleven full cards asked, ten there: the message quotes all twelve cells and the ten indexes placed so far
    std::vector<int> eleven(11, 100);
    auto f = d.Assume({"wide"}, MockPodWithDemand(eleven), PolicySpec{}, false);
    CHECK(!f.first[0]);
    CHECK_EQ_STR(f.second[0],
                 "can't allocate   [(100) (100) (100) (100) (100) (100) (100) (100) (100) (100) (100)] on "
                 "(100)(100)(100)(0)(100)(100)(100)(100)(100)(100)(40)(100): indexes=[0 1 2 5 6 7 8 9 4 11]");
    Pod p = MockPodWithDemand({50, 50, 30}, "p1");
    p.node_name = "wide";
    CHECK(d.Bind("wide", &p, PolicySpec{}, false).empty());
    CHECK((plan_of(p) == std::vector<int>{0, 0, 10}));
    std::vector<int> after = start;
    after[0] = 0;
    after[10] = 10;
    CHECK((d.Status()["wide"] == after));
    Pod big = MockPodWithPlan({20}, {10}, "wide", "big");
    CHECK_EQ_STR(d.Allocate(big), "can't apply plan &{[(20)] [10] 0} on (0)(100)(100)(0)(100)(100)(100)(100)(100)(100)(10)(100)");  // allocate.go:113
    Pod fits = MockPodWithPlan({70}, {11}, "wide", "fits");
    CHECK(d.Allocate(fits).empty());
    after[11] = 30;
    CHECK((d.Status()["wide"] == after));
    CHECK(d.Release(p).empty());
    after[0] = 100;
    after[10] = 40;
    CHECK((d.Status()["wide"] == after));
    CHECK(d.StatusJSON().find("\"wide\":{\"Rater\":{},\"Name\":\"wide\",\"GPUs\":[{\"Percent\":100,") != std::string::npos);
    CHECK(d.StatusJSON().find("{\"Percent\":40,\"PercentTotal\":100,\"RemainLoad\":0},{\"Percent\":30,\"PercentTotal\":100,\"RemainLoad\":0}],") != std::string::npos);
}

static void TestLoadAwareScoring() {
    FakeCluster fc;
    fc.add("busy", 2);
    fc.add("idle", 2);
    Dealer d(&fc, RaterKind::Binpack, 16);
    PolicySpec spec;
    spec.sync_period = {{"gpu_core_usage_avg", 15LL * 1000000000LL}, {"gpu_memory_usage_avg", 15LL * 1000000000LL}};
    // 2024-03-01T12:00:10Z in Asia/Shanghai
    d.SetNow(1709265610LL * 1000000000LL);
    for (int card = 0; card < 2; card++) {
        d.AddCoreUsage("busy");  // (re-)creating the map drops earlier cards, as nodeusage.go:61 does
        d.AddMemoryUsage("busy");
    }
    for (int card = 0; card < 2; card++) {
        d.UpdateCoreUsage("busy", "0.95000", "2024-03-01T12:00:00Z", card);
        d.UpdateMemoryUsage("busy", "0.85000", "2024-03-01T12:00:00Z", card);
    }
    auto u = d.GetUsage("busy", "gpu_core_usage_avg", 0, 315LL * 1000000000LL);
    CHECK(u.exist && u.err.empty() && u.value == 0.95);
    CHECK(!d.GetUsage("idle", "gpu_core_usage_avg", 0, 315LL * 1000000000LL).exist);
    CHECK_EQ_STR(d.GetUsage("busy", "gpu_memory_usage_avg", 1, 5).err, "gpu_memory_usage_avgnot in update period");  // nodeusage.go:100
    Pod pod = MockPodWithDemand({30});
    std::vector<int> plain = d.Score({"busy", "idle"}, pod, spec, false);
    CHECK(plain[0] == -2 && plain[1] == -2);  // fresh 2-GPU nodes: 0 + 0 - 2 (rater.go:69)
    auto a = d.Assume({"busy", "idle"}, MockPodWithDemand({31}), spec, true);
    CHECK(a.first[0] && a.first[1]);
    std::vector<int> sc = d.Score({"busy", "idle"}, MockPodWithDemand({31}), spec, true);
    // loadUsage(busy) = 2 cards x (ceil(9.5)/10 + ceil(8.5)/10) = 3.8 -> int 3 / 2 GPUs = 1 -> +50 (rater.go:68-69)
    CHECK(sc[0] == 48 && sc[1] == -2);
}

// the controller's metric sync (controller/node.go:111-135, prometheus.go:57-62) feeding the same tables
static void TestAnnotatorNodeIngestion() {
    CHECK_EQ_STR(FormatLocalTime(1709265610LL * 1000000000LL), "2024-03-01T12:00:10Z");  // 04:00:10 UTC shown as Shanghai wall clock
    CHECK_EQ_STR(FormatLocalTime(1709251200LL * 1000000000LL + 999999999LL), "2024-03-01T08:00:00Z");
    CHECK_EQ_STR(FormatMetricValue(0.949996), "0.95000");
    CHECK_EQ_STR(FormatMetricValue(-0.3), "0.00000");
    CHECK_EQ_STR(FormatMetricValue(0.0 / 0.0), "0.00000");
    int64_t back = 0;
    CHECK(ngs_parse_update_time(FormatLocalTime(1709265610LL * 1000000000LL).c_str(), &back) == 1 && back == 1709265610LL);
    FakeCluster fc;
    fc.add("busy", 2);
    fc.add("idle", 2);
    Dealer d(&fc, RaterKind::Binpack, 16), ref(&fc, RaterKind::Binpack, 16);
    PolicySpec spec;
    spec.sync_period = {{"gpu_core_usage_avg", 15LL * 1000000000LL}, {"gpu_memory_usage_avg", 15LL * 1000000000LL}};
    d.SetNow(1709265600LL * 1000000000LL);
    ref.SetNow(1709265600LL * 1000000000LL);
    for (int card = 0; card < 2; card++) {
        CHECK(d.AnnotatorNode("busy", "gpu_core_usage_avg", std::to_string(card), FormatMetricValue(0.95)).empty());
        CHECK(d.AnnotatorNode("busy", "gpu_memory_usage_avg", std::to_string(card), FormatMetricValue(0.85)).empty());
    }
    CHECK(d.GetCoreUsageLock("busy").second && d.GetCoreUsage("busy").first.size() == 2);  // the map was created once, both cards kept
    CHECK_EQ_STR(d.GetCoreUsage("busy").first[1].UpdateTime, "2024-03-01T12:00:00Z");
    CHECK_EQ_STR(d.AnnotatorNode("busy", "gpu_core_usage_avg", "0", ""),
                 "QueryLasterData gpu_core_usage_avg for node busy return value:  , error: <nil>");
    CHECK(!d.AnnotatorNode("busy", "gpu_core_usage_avg", "x", "0.5").empty());
    ref.AddCoreUsage("busy");
    ref.AddMemoryUsage("busy");
    for (int card = 0; card < 2; card++) {
        ref.UpdateCoreUsage("busy", "0.95000", "2024-03-01T12:00:00Z", card);
        ref.UpdateMemoryUsage("busy", "0.85000", "2024-03-01T12:00:00Z", card);
    }
    d.SetNow(1709265610LL * 1000000000LL);  // ten seconds later the filter runs
    ref.SetNow(1709265610LL * 1000000000LL);
    Pod pod = MockPodWithDemand({31});
    CHECK(d.Assume({"busy", "idle"}, pod, spec, true).first[0]);
    CHECK(ref.Assume({"busy", "idle"}, pod, spec, true).first[0]);
    CHECK(d.Score({"busy", "idle"}, pod, spec, true) == ref.Score({"busy", "idle"}, pod, spec, true));
    CHECK(d.Score({"busy", "idle"}, pod, spec, true)[0] == 48);
}

// the informer side: syncPod / updatePodInCache / deletePodFromCache (controller.go:210-357) drive Allocate, Release, Forget
static void TestControllerPodEvents() {
    FakeCluster fc;
    fc.add("n", 2);
    Dealer d(&fc, RaterKind::Binpack, 8);
    Pod pending = MockPodWithDemand({60, 30}, "p1");
    CHECK(SyncPod(d, nullptr).forget);                                    // deleted before the worker got to it
    CHECK(SyncPod(d, &pending).forget && d.Status().empty());             // not scheduled yet: nothing to account
    Pod running = MockPodWithPlan({60, 30}, {1, 1}, "n", "p1");           // the scheduler bound it; the informer sees the update
    CHECK(NeedUpdate(d, pending, running));                               // unknown, unreleased, annotation set
    SyncResult r = SyncPod(d, &running);
    CHECK(r.forget && r.err.empty() && d.KnownPod(running));
    CHECK((d.Status()["n"] == std::vector<int>{100, 10}));
    CHECK(!NeedUpdate(d, running, running));                              // known and still running: no work
    Pod bad = MockPodWithPlan({60}, {1}, "n", "p2");                      // a plan that no longer fits
    r = SyncPod(d, &bad);
    CHECK(!r.forget && !r.err.empty());                                   // re-queued with the Allocate error (:238-240)
    Pod done = running;
    done.completed = true;
    CHECK(NeedUpdate(d, running, done));                                  // known pod turned complete
    CHECK(SyncPod(d, &done).forget && d.PodReleased(done) && !d.KnownPod(done));
    CHECK((d.Status()["n"] == std::vector<int>{100, 100}));
    CHECK(!NeedUpdate(d, done, done));                                    // released: later updates are ignored
    DeletePodFromCache(d, done);
    CHECK(!d.PodReleased(done));                                          // Forget drops the bookkeeping, no capacity moves
    CHECK((d.Status()["n"] == std::vector<int>{100, 100}));
}

// DealerImpl serves the extender's HTTP handlers concurrently behind d.Lock (dealer.go:81): four
// threads filter, score, bind and release against one Dealer; afterwards no capacity is lost or made up.
static void TestConcurrentCallers() {
    FakeCluster fc;
    std::vector<std::string> names;
    for (int i = 0; i < 24; i++) {
        fc.add("node-" + std::to_string(i), 4);
        names.push_back("node-" + std::to_string(i));
    }
    Dealer d(&fc, RaterKind::Spread, 64);
    std::atomic<int> bound_total{0}, released_total{0}, errors{0};
    auto worker = [&](int t) {
        std::vector<Pod> mine;
        for (int k = 0; k < 60; k++) {
            Pod pod = MockPodWithDemand({20 + 5 * (k % 3), 10}, "t" + std::to_string(t) + "-" + std::to_string(k));
            auto r = d.Assume(names, pod, PolicySpec{}, false);
            std::vector<int> sc = d.Score(names, pod, PolicySpec{}, false);
            int best = -1;
            for (size_t i = 0; i < names.size(); i++)
                if (r.first[i] && (best < 0 || sc[i] > sc[best])) best = (int)i;
            if (best < 0) continue;
            pod.node_name = names[best];
            // another thread may have taken the capacity between Assume and Bind: an error here is legal, corruption is not
            if (d.Bind(names[best], &pod, PolicySpec{}, false).empty()) {
                mine.push_back(pod);
                bound_total++;
            }
            if (k % 4 == 3 && !mine.empty()) {
                if (!d.Release(mine.back()).empty()) errors++;
                mine.pop_back();
                released_total++;
            }
        }
        for (Pod& p : mine) {
            if (!d.Release(p).empty()) errors++;
            released_total++;
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < 4; t++) th.emplace_back(worker, t);
    for (auto& x : th) x.join();
    CHECK(errors == 0);
    CHECK(bound_total > 100 && bound_total == released_total);
    auto st = d.Status();
    CHECK(st.size() == names.size());
    for (auto& kv : st)
        for (int f : kv.second) CHECK(f == 100);  // everything that was bound came back, cell by cell
}

int main() {
    {
        FakeCluster fc;
        Dealer probe(&fc, RaterKind::Binpack, 4);
        if (!probe.ok()) {
            printf("no usable CUDA device: %s\n", probe.init_error().c_str());
            return 2;
        }
    }
    TestBinpackChoose();
    TestSpreadChoose();
    TestChoose();
    TestBinpackRate();
    TestSpreadRate();
    TestNewDemandAndPlanFromPod();
    TestAssumeErrorTexts();
    TestBindCommitFailureLeavesPodUnknown();
    ScheduleCycleAndRelease(false);
    ScheduleCycleAndRelease(true);
    TestAllocateFromAnnotationsAndRebuild();
    TestTwelveGpuNode();
    TestLoadAwareScoring();
    TestAnnotatorNodeIngestion();
    TestControllerPodEvents();
    TestConcurrentCallers();
    printf("%s: %d checks, %d failed\n", g_fail ? "HOST TESTS FAILED" : "HOST TESTS OK", g_checks, g_fail);
    return g_fail ? 1 : 0;
}
