// UNVERIFIED: written without a Go toolchain (none in the build environment); never compiled.

//go:build ksareplay
// +build ksareplay

package replay

import (
	"testing"
	"time"
)

func maskOf(t *testing.T, names []string) uint32 {
	var mask uint32
	for _, n := range names {
		m, err := ParseAttributeNames(n)
		if err != nil {
			t.Fatal(err)
		}
		mask |= m
	}
	return mask
}

// TestGPUMatchesReference runs every fixture through the reference (runVirtual, reference_test.go)
// AND through the GPU library (Pipeline.Replay, host slices in and out) and compares the two
// directly, tick by tick: the comparison north_star asks for, with nothing of this repository's
// oracle in between. Needs a CUDA device and libksareplay (build tag ksareplay).
func TestGPUMatchesReference(t *testing.T) {
	for _, c := range loadFixtures(t).Cases {
		c := c
		t.Run(c.ID, func(t *testing.T) {
			wantReplicas, wantEvents := runVirtual(t, c)
			pol := Policy{
				PollInterval:        time.Duration(c.Policy.PollPeriodNs),
				ScaleDownCoolPeriod: time.Duration(c.Policy.ScaleDownCoolDownNs),
				ScaleUpCoolPeriod:   time.Duration(c.Policy.ScaleUpCoolDownNs),
				ScaleUpMessages:     c.Policy.ScaleUpMessages, ScaleDownMessages: c.Policy.ScaleDownMessages,
				ScaleUpPods: c.Policy.ScaleUpPods, ScaleDownPods: c.Policy.ScaleDownPods,
				MaxPods: c.Policy.MaxPods, MinPods: c.Policy.MinPods, AttrMask: maskOf(t, c.Policy.AttributeNames),
			}
			T, A := len(c.Depth), pol.Attrs()
			depth := make([]int64, 0, T*A)
			for _, row := range c.Depth { // [T][A][N] with N = 1
				depth = append(depth, row...)
			}
			pipe, err := NewPipeline(0, 1, A, 16)
			if err != nil {
				t.Fatal(err)
			}
			defer pipe.Close()
			st := NewState(1, c.Replicas0)
			replicas, events := make([]int32, T), make([]uint8, T)
			if err := pipe.Replay(pol, depth, c.Fail, T, 0, st, replicas, events, nil); err != nil {
				t.Fatal(err)
			}
			for k := 0; k < T; k++ {
				if replicas[k] != wantReplicas[k] || events[k] != wantEvents[k] {
					t.Fatalf("%s: tick %d: GPU has replicas %d, events %#x; the reference has %d, %#x",
						c.ID, k+1, replicas[k], events[k], wantReplicas[k], wantEvents[k])
				}
			}
		})
	}
}
