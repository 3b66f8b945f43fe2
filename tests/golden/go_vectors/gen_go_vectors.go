// gen_go_vectors.go — prints JSON vectors from the REFERENCE's own pkg/dealer (Go 1.16, go.mod:3) for exactly the three
// areas of the oracle no reference test pins (DESIGN.md "Oracle pinning status"):
//   1. sort.Sort tie order at 7..12 GPUs through Binpack/Spread.Choose (rater.go:74-163, allocate.go:243-247)
//   2. Binpack.Rate's fp64 truncation over every used total of 1..12 GPUs (rater.go:59-70, allocate.go:164-171)
//   3. the load-aware mode: Rate, the RemainLoad it leaves (allocate.go:173-195) and Choose on seeded rows, with
//      DealerImpl.CoreUsage / MemoryUsage pre-populated (dealer.go:84-85, nodeusage.go:82-111, stats.go:30-66)
// It cannot run in the build container of this repository (no Go toolchain, un-vendored k8s.io modules).  Where Go 1.16
// and the reference's module cache exist:
//     cp tests/golden/go_vectors/gen_go_vectors.go <reference>/cmd/genvectors/main.go
//     cd <reference> && go run ./cmd/genvectors > <this repo>/tests/golden/go_vectors.json
// tests/test_oracle_golden.py::test_go_vectors_when_present then checks the oracle against every vector.
// Timestamps are written relative to the generator's time.Now(): each usage sample carries "age_s" (seconds before now);
// the consumer rebuilds absolute times around its own fixed clock, so the file stays valid forever.
package main

import (
	"encoding/json"
	"fmt"
	"math/rand"
	"os"
	"time"

	"github.com/nano-gpu/nano-gpu-scheduler/pkg/dealer"
)

type chooseVec struct {
	Policy string `json:"policy"`
	Free   []int  `json:"free"`
	RL     []int  `json:"remain_load"`
	Demand []int  `json:"demand"`
	OK     bool   `json:"ok"`
	Plan   []int  `json:"plan"`
}
type rateVec struct {
	G     int `json:"gpus"`
	Used  int `json:"used"`
	Score int `json:"score"`
}
type sample struct {
	Value string `json:"value"` // as the metric sync writes it (%.5f), or garbage
	AgeS  int    `json:"age_s"` // update time = now - age_s
	Bad   bool   `json:"bad_time"`
}
type loadVec struct {
	Policy     string     `json:"policy"`
	Free       []int      `json:"free"`
	Periods    [][2]int64 `json:"periods"` // (metric: 0 core / 1 memory, period in ns) in SyncPeriod order
	Core       []sample   `json:"core"`    // nil: the node has no entry in the table
	Memory     []sample   `json:"memory"`
	Demand     []int      `json:"demand"`
	Score      int        `json:"score"`
	RemainLoad []int      `json:"remain_load_after"`
	OK         bool       `json:"ok"`
	Plan       []int      `json:"plan"`
}

func gpus(free, rl []int) dealer.GPUs {
	g := dealer.GPUs{}
	for i := range free {
		r := &dealer.GPUResource{Percent: free[i], PercentTotal: 100}
		if rl != nil {
			r.RemainLoad = rl[i]
		}
		g = append(g, r)
	}
	return g
}
func demand(d []int) dealer.Demand {
	out := dealer.Demand{}
	for _, p := range d {
		out = append(out, dealer.GPUResource{Percent: p})
	}
	return out
}
func rater(policy string) dealer.Rater {
	if policy == "spread" {
		return &dealer.Spread{}
	}
	return &dealer.Binpack{}
}

func main() {
	rng := rand.New(rand.NewSource(0xB200))
	out := map[string]interface{}{"go": "1.16", "generated": time.Now().UTC().Format(time.RFC3339)}

	// 1. tie order: many equal keys, 7..12 GPUs, RemainLoad mixed in (key = free + 50*RemainLoad)
	var cv []chooseVec
	for _, policy := range []string{"binpack", "spread"} {
		for n := 0; n < 400; n++ {
			G := 7 + rng.Intn(6)
			free, rl := make([]int, G), make([]int, G)
			for i := range free {
				free[i] = 20 * rng.Intn(6) // few distinct values: ties everywhere
				if n%3 == 0 {
					rl[i] = rng.Intn(3)
				}
			}
			C := 1 + rng.Intn(5)
			d := make([]int, C)
			for i := range d {
				d[i] = 10 * rng.Intn(7)
			}
			plan, err := rater(policy).Choose(gpus(free, rl), demand(d))
			cv = append(cv, chooseVec{policy, free, rl, d, err == nil, plan})
		}
	}
	out["choose"] = cv

	// 2. Binpack.Rate over every total
	var rv []rateVec
	for G := 1; G <= 12; G++ {
		for used := 0; used <= 100*G; used++ {
			free := make([]int, G)
			left := used
			for i := range free {
				take := left
				if take > 100 {
					take = 100
				}
				free[i] = 100 - take
				left -= take
			}
			s := (&dealer.Binpack{}).Rate(gpus(free, nil), nil, nil, dealer.PolicySpec{}, "n", false)
			rv = append(rv, rateVec{G, used, s})
		}
	}
	out["binpack_rate"] = rv

	// 3. load-aware mode
	loc, err := time.LoadLocation("Asia/Shanghai")
	if err != nil {
		loc = time.FixedZone("CST", 8*3600)
	}
	stamp := func(age int) string { return time.Now().In(loc).Add(-time.Duration(age) * time.Second).Format("2006-01-02T15:04:05Z") }
	var lv []loadVec
	for n := 0; n < 600; n++ {
		policy := []string{"binpack", "spread"}[n%2]
		G := 1 + rng.Intn(8)
		free := make([]int, G)
		for i := range free {
			free[i] = 5 * rng.Intn(21)
		}
		spec := dealer.PolicySpec{}
		var periods [][2]int64
		for _, m := range [][]int{{0, 1}, {1, 0}, {0}, {1}, {0, 0, 1}, {}}[n%6] {
			name := dealer.GPUCoreUsagePriority
			if m == 1 {
				name = dealer.GPUMemoryUsagePriority
			}
			p := time.Duration(15+rng.Intn(3)*15) * time.Second
			spec.SyncPeriod = append(spec.SyncPeriod, dealer.Period{Name: name, Period: p})
			periods = append(periods, [2]int64{int64(m), int64(p)})
		}
		mk := func() []sample {
			if rng.Intn(12) == 0 {
				return nil
			}
			s := make([]sample, G)
			for i := range s {
				v := fmt.Sprintf("%.5f", float64(rng.Intn(100001))/100000.0)
				switch rng.Intn(15) {
				case 0:
					v = "1.20000"
				case 1:
					v = "abc"
				}
				age := rng.Intn(200)
				if rng.Intn(10) == 0 {
					age = 100000 + rng.Intn(1000) // stale
				}
				s[i] = sample{v, age, rng.Intn(25) == 0}
			}
			return s
		}
		core, mem := mk(), mk()
		d := &dealer.DealerImpl{CoreUsage: map[string]map[int]dealer.GPUCoreUsage{}, MemoryUsage: map[string]map[int]dealer.GPUMemoryUsage{}}
		if core != nil {
			d.CoreUsage["n"] = map[int]dealer.GPUCoreUsage{}
			for i, s := range core {
				t := stamp(s.AgeS)
				if s.Bad {
					t = "not-a-time"
				}
				d.CoreUsage["n"][i] = dealer.NewGPUCoreUsage(s.Value, t)
			}
		}
		if mem != nil {
			d.MemoryUsage["n"] = map[int]dealer.GPUMemoryUsage{}
			for i, s := range mem {
				t := stamp(s.AgeS)
				if s.Bad {
					t = "not-a-time"
				}
				d.MemoryUsage["n"][i] = dealer.NewGPUMemoryUsage(s.Value, t)
			}
		}
		C := 1 + rng.Intn(4)
		dm := make([]int, C)
		for i := range dm {
			dm[i] = 5 * rng.Intn(13)
		}
		g := gpus(free, nil)
		plan, err := g.Choose(demand(dm), rater(policy), d, spec, "n", true) // Rate (writes RemainLoad), then Choose
		rl := make([]int, G)
		for i := range g {
			rl[i] = g[i].RemainLoad
		}
		lv = append(lv, loadVec{policy, free, periods, core, mem, dm, plan.Score, rl, err == nil, plan.GPUIndexes})
	}
	out["load"] = lv
	enc := json.NewEncoder(os.Stdout)
	enc.SetIndent("", " ")
	if err := enc.Encode(out); err != nil {
		panic(err)
	}
}
