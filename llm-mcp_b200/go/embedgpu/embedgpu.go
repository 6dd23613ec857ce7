// Package embedgpu binds libembed_truncate_norm.so (include/embed_truncate_norm.h) for
// LLM-MCP's core.  It is the cgo package INTEGRATION.md proposes as
// core/internal/embedgpu; TruncateEmbeddings is a drop-in for the private function
// truncateEmbeddings(result map[string]any, dims int) at core/internal/api/handlers.go:2062-2078.
//
// NOT COMPILED IN THIS REPOSITORY: the build image has no Go toolchain (SURVEY.md F4).
// The Python mirror llm-mcp_b200/embedgpu.py implements the same logic over the same C-ABI
// and is what the tests exercise.  Building this needs CGO_ENABLED=1 (the reference builds
// with CGO_ENABLED=0, core/Dockerfile:6) and a glibc base image with the NVIDIA driver.
package embedgpu

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../.. -lembed_truncate_norm
#include <stdlib.h>
#include "embed_truncate_norm.h"
*/
import "C"

import (
	"fmt"
	"log/slog"
	"os"
	"strconv"
	"sync/atomic"
	"unsafe"
)

// Init binds the visible GPUs (call once from main, next to api.New at
// core/cmd/core/main.go:49-51).  A non-nil error means "serve with the reference reslice".
func Init() error {
	if rc := C.etn_init(nil, 0); rc != 0 {
		return fmt.Errorf("etn_init: %s (%d)", C.GoString(C.etn_strerror(rc)), int(rc))
	}
	ready.Store(true)
	return nil
}

// Shutdown releases workers, streams and staging buffers (main.go:118-122) and the pinned
// buffers parked in the free list.  Handlers still inside a call keep their buffers until they
// return; those are then freed by put() because the list is closed.
func Shutdown() {
	ready.Store(false)
	for {
		select {
		case p := <-free:
			p.release()
		default:
			C.etn_shutdown()
			return
		}
	}
}

// ready is read by every handler goroutine and written by Init/Shutdown.
var ready atomic.Bool

// RenormEnabled reports EMBED_RENORMALIZE=1, read on every call like config.GetenvInt does
// (core/internal/config/config.go:17-24).  Default off: the reference only truncates.
func RenormEnabled() bool {
	n, err := strconv.Atoi(os.Getenv("EMBED_RENORMALIZE"))
	return err == nil && n == 1
}

// pinned is a reusable pair of page-locked buffers; handlers run concurrently
// (one goroutine per connection), so they come from a pool.
type pinned struct {
	in, out   unsafe.Pointer
	inN, outN int // capacity in float64 elements
}

// free is a bounded free list.  A sync.Pool would be wrong here: the GC empties it every
// couple of cycles and the page-locked memory behind a dropped *pinned (cudaHostAlloc, ~40 MiB
// for a C1-sized response) would never be returned.  At most cap(free) buffer pairs stay
// parked; a pair that does not fit is released on the spot.
var free = make(chan *pinned, 16)

func get() *pinned {
	select {
	case p := <-free:
		return p
	default:
		return &pinned{}
	}
}

func put(p *pinned) {
	if ready.Load() {
		select {
		case free <- p:
			return
		default:
		}
	}
	p.release()
}

func (p *pinned) release() {
	C.etn_free_pinned(p.in)
	C.etn_free_pinned(p.out)
	p.in, p.out, p.inN, p.outN = nil, nil, 0, 0
}

// ensure grows the buffers; a capacity is recorded only for an allocation that succeeded, so a
// failed one is retried by the next request instead of poisoning the pair.
func (p *pinned) ensure(inN, outN int) bool {
	if inN > p.inN {
		C.etn_free_pinned(p.in)
		p.in, p.inN = C.etn_alloc_pinned(C.size_t(inN*8)), 0
		if p.in != nil {
			p.inN = inN
		}
	}
	if outN > p.outN {
		C.etn_free_pinned(p.out)
		p.out, p.outN = C.etn_alloc_pinned(C.size_t(outN*8)), 0
		if p.out != nil {
			p.outN = outN
		}
	}
	return p.in != nil && p.out != nil
}

// reslice is the reference behaviour, kept as the fallback (handlers.go:2073-2076).
func reslice(m map[string]any, emb []any, dims int) { m["embedding"] = emb[:dims] }

// TruncateEmbeddings keeps the first dims components of every embedding of an OpenAI-format
// response and, when EMBED_RENORMALIZE=1, L2-renormalises them on the GPU.  Same skips as
// the reference: data not a list, items not maps, embeddings that are not lists (base64,
// null), vectors with len <= dims.  Vectors are packed per source length into pinned
// memory (no Go pointer is ever handed to C) and each group is one
// embed_truncate_norm_host call.  Any error falls back to the reslice, so the request
// still succeeds (the cloud branch has no retry, handlers.go:1881-1886).
func TruncateEmbeddings(result map[string]any, dims int) {
	data, ok := result["data"].([]any)
	if !ok || dims <= 0 {
		return
	}
	mode := C.int(C.ETN_TRUNCATE)
	if RenormEnabled() {
		mode = C.ETN_TRUNCATE_RENORM
	}
	groups := map[int][]map[string]any{}
	for _, item := range data {
		m, ok := item.(map[string]any)
		if !ok {
			continue
		}
		emb, ok := m["embedding"].([]any)
		if !ok || len(emb) <= dims {
			continue
		}
		if !ready.Load() || mode == C.ETN_TRUNCATE || !allFloat64(emb) {
			reslice(m, emb, dims) // truncation alone needs no arithmetic: O(1), as in the reference
			if mode != C.ETN_TRUNCATE {
				C.etn_note_fallback(1) // renormalisation was asked for and could not be done on the GPU
			}
			continue
		}
		groups[len(emb)] = append(groups[len(emb)], m)
	}
	for D, items := range groups {
		n := len(items)
		p := get()
		if !p.ensure(n*D, n*dims) {
			put(p)
			for _, m := range items {
				reslice(m, m["embedding"].([]any), dims)
			}
			continue
		}
		in := unsafe.Slice((*float64)(p.in), n*D)
		for r, m := range items {
			for c, v := range m["embedding"].([]any) {
				in[r*D+c] = v.(float64)
			}
		}
		rc := C.embed_truncate_norm_host(C.int64_t(n), C.int64_t(D), C.int64_t(dims), C.ETN_F64, mode,
			p.in, 0, p.out, 0)
		if rc != 0 {
			slog.Error("embedgpu", "error", C.GoString(C.etn_strerror(rc)), "code", int(rc))
			for _, m := range items {
				reslice(m, m["embedding"].([]any), dims)
			}
			put(p)
			continue
		}
		out := unsafe.Slice((*float64)(p.out), n*dims)
		for r, m := range items {
			emb := m["embedding"].([]any)[:dims] // reuse the backing array, as the reslice does
			for c := range emb {
				emb[c] = out[r*dims+c]
			}
			m["embedding"] = emb
		}
		put(p)
	}
}

func allFloat64(emb []any) bool {
	for _, v := range emb {
		if _, ok := v.(float64); !ok {
			return false
		}
	}
	return true
}

// TruncateTyped is the op for the local Ollama branch (SURVEY.md §8(f) N2), where vectors
// arrive typed: Embeddings [][]float64 (handlers.go:1976-1984) and `dimensions` is ignored
// today.  Same rules as TruncateEmbeddings; without EMBED_RENORMALIZE it is a reslice.  With
// it, the vectors are grouped by source length and every group is ONE host call (as the
// Python mirror does; a call per vector would pay the launch chain once per vector).
func TruncateTyped(embs [][]float64, dims int) {
	if dims <= 0 {
		return
	}
	renorm := ready.Load() && RenormEnabled()
	groups := map[int][]int{}
	for i, e := range embs {
		if len(e) <= dims {
			continue
		}
		if !renorm {
			embs[i] = e[:dims]
			continue
		}
		groups[len(e)] = append(groups[len(e)], i)
	}
	for D, idx := range groups {
		n := len(idx)
		p := get()
		ok := p.ensure(n*D, n*dims)
		if ok {
			in := unsafe.Slice((*float64)(p.in), n*D)
			for r, i := range idx {
				copy(in[r*D:(r+1)*D], embs[i])
			}
			rc := C.embed_truncate_norm_host(C.int64_t(n), C.int64_t(D), C.int64_t(dims), C.ETN_F64, C.ETN_TRUNCATE_RENORM,
				p.in, 0, p.out, 0)
			if rc != 0 {
				slog.Error("embedgpu", "error", C.GoString(C.etn_strerror(rc)), "code", int(rc))
				ok = false
			}
		}
		if ok {
			out := unsafe.Slice((*float64)(p.out), n*dims)
			for r, i := range idx {
				copy(embs[i][:dims], out[r*dims:(r+1)*dims])
			}
		} else {
			C.etn_note_fallback(C.int64_t(n))
		}
		for _, i := range idx {
			embs[i] = embs[i][:dims]
		}
		put(p)
	}
}

// Counters returns the library's cumulative counters keyed by Prometheus metric name
// (SURVEY.md §8(f) N3); register them as CounterFuncs next to metrics.go:9-26.
func Counters() map[string]int64 {
	var buf [C.ETN_CTR_COUNT]C.int64_t
	n := int(C.etn_get_counters(&buf[0], C.ETN_CTR_COUNT))
	out := make(map[string]int64, n)
	for i := 0; i < n; i++ {
		out[C.GoString(C.etn_counter_name(C.int(i)))] = int64(buf[i])
	}
	return out
}

// Histogram is one of the library's fixed-bucket histograms (ETN_HIST_*): register it beside
// llmcore_embedding_duration_seconds (metrics.go:18-26) through a prometheus.Collector that
// emits prometheus.MustNewConstHistogram(desc, Count, Sum, Buckets).
type Histogram struct {
	Name, Help string
	Buckets    map[float64]uint64 // cumulative, keyed by upper bound (without +Inf), as client_golang wants
	Count      uint64
	Sum        float64
}

// Histograms reads all of them.
func Histograms() []Histogram {
	out := make([]Histogram, 0, int(C.ETN_HIST_COUNT))
	for w := 0; w < int(C.ETN_HIST_COUNT); w++ {
		var le [32]C.double
		var cnt [32]C.int64_t
		var sum C.double
		n := int(C.etn_get_histogram(C.int(w), &le[0], &cnt[0], 32, &sum))
		if n <= 0 {
			continue
		}
		h := Histogram{Name: C.GoString(C.etn_histogram_name(C.int(w))), Help: C.GoString(C.etn_histogram_help(C.int(w))),
			Buckets: map[float64]uint64{}, Sum: float64(sum)}
		var run uint64
		for i := 0; i < n; i++ {
			run += uint64(cnt[i])
			if i < n-1 {
				h.Buckets[float64(le[i])] = run
			}
		}
		h.Count = run
		out = append(out, h)
	}
	return out
}
