// Go benchmark of the reference path for anyone who has a Go toolchain (this image has none:
// `go version` fails here and on the GPU box, so bench.py's reference arm times the C
// restatement oracle/etn_oracle.c instead; SURVEY.md F4).
//
//	cd baseline/go && go mod init etnbench && go test -bench . -benchmem
//
// BenchmarkReslice is what LLM-MCP does today (truncateEmbeddings,
// core/internal/api/handlers.go:2062-2078): an O(1) slice-header update per vector.
// BenchmarkTruncateRenorm is the loop a Go author would add for the renormalise step that
// BASELINE.json defines; it is the loop oracle O-go models (sequential float64, no FMA at
// GOAMD64=v1).  Shape: C1, 1,024 vectors of 4096 float64 cut to 1024.
package etnbench

import (
	"math"
	"runtime"
	"sync"
	"testing"
)

const (
	batch = 1024
	dim   = 4096
	keep  = 1024
)

func response() map[string]any {
	data := make([]any, batch)
	for i := range data {
		emb := make([]any, dim)
		for j := range emb {
			emb[j] = float64((i*dim+j)%2001-1000) / 32000.0
		}
		data[i] = map[string]any{"object": "embedding", "index": float64(i), "embedding": emb}
	}
	return map[string]any{"object": "list", "data": data}
}

// reslice restates the reference's behaviour (handlers.go:2062-2078) for the benchmark: for
// every object in result["data"] whose "embedding" is a []any longer than dims, keep the
// first dims entries by re-slicing; nothing else is touched.
func reslice(result map[string]any, dims int) (cut int) {
	items, _ := result["data"].([]any)
	for i := range items {
		if obj, isObj := items[i].(map[string]any); isObj {
			if vec, isVec := obj["embedding"].([]any); isVec && len(vec) > dims {
				obj["embedding"] = vec[:dims]
				cut++
			}
		}
	}
	return cut
}

func BenchmarkReslice(b *testing.B) {
	r := response()
	full := make([][]any, batch)
	for i, it := range r["data"].([]any) {
		full[i] = it.(map[string]any)["embedding"].([]any)
	}
	b.ResetTimer()
	for n := 0; n < b.N; n++ {
		for i, it := range r["data"].([]any) { // restore, then cut
			it.(map[string]any)["embedding"] = full[i]
		}
		reslice(r, keep)
	}
	b.ReportMetric(float64(batch)*float64(b.N)/b.Elapsed().Seconds(), "vectors/s")
}

func renormRow(e []float64, out []float64) {
	ss := 0.0
	for _, v := range e {
		ss += v * v
	}
	n := math.Sqrt(ss)
	if n > 0 {
		for i, v := range e {
			out[i] = v / n
		}
	} else {
		copy(out, e)
	}
}

func BenchmarkTruncateRenorm(b *testing.B) {
	x := make([]float64, batch*dim)
	for i := range x {
		x[i] = float64(i%2001-1000) / 32000.0
	}
	y := make([]float64, batch*keep)
	workers := runtime.GOMAXPROCS(0)
	b.ResetTimer()
	for n := 0; n < b.N; n++ {
		var wg sync.WaitGroup
		for w := 0; w < workers; w++ {
			wg.Add(1)
			go func(w int) {
				defer wg.Done()
				for r := w; r < batch; r += workers {
					renormRow(x[r*dim:r*dim+keep], y[r*keep:(r+1)*keep])
				}
			}(w)
		}
		wg.Wait()
	}
	b.ReportMetric(float64(batch)*float64(b.N)/b.Elapsed().Seconds(), "vectors/s")
	b.ReportMetric(float64(workers), "GOMAXPROCS")
}
