// Tests in the style of core/internal/api/helpers_test.go.  NOT RUN HERE (no Go toolchain);
// tests/test_gpu_host.py holds the same cases against the Python mirror.
package embedgpu

import (
	"math"
	"testing"
)

func resp(vs ...[]any) map[string]any {
	data := make([]any, len(vs))
	for i, v := range vs {
		data[i] = map[string]any{"object": "embedding", "index": float64(i), "embedding": v}
	}
	return map[string]any{"object": "list", "data": data}
}

func TestTruncateMatchesReference(t *testing.T) {
	r := resp([]any{3.0, 4.0, 0.0, 0.0}, []any{1.0, 2.0}, []any{1.0, "x", 3.0})
	TruncateEmbeddings(r, 2)
	d := r["data"].([]any)
	if e := d[0].(map[string]any)["embedding"].([]any); len(e) != 2 || e[0] != 3.0 || e[1] != 4.0 {
		t.Fatalf("row 0: %v", e)
	}
	if e := d[1].(map[string]any)["embedding"].([]any); len(e) != 2 {
		t.Fatalf("row 1 must be left whole: %v", e)
	}
	if e := d[2].(map[string]any)["embedding"].([]any); len(e) != 2 || e[1] != "x" {
		t.Fatalf("row 2 must be resliced: %v", e)
	}
}

func TestRenorm(t *testing.T) {
	if err := Init(); err != nil {
		t.Skip(err)
	}
	defer Shutdown()
	t.Setenv("EMBED_RENORMALIZE", "1")
	r := resp([]any{3.0, 4.0, 0.0, 0.0})
	TruncateEmbeddings(r, 2)
	e := r["data"].([]any)[0].(map[string]any)["embedding"].([]any)
	if math.Abs(e[0].(float64)-0.6) > 1e-15 || math.Abs(e[1].(float64)-0.8) > 1e-15 {
		t.Fatalf("got %v", e)
	}
}
