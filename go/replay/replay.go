// UNVERIFIED: written without a Go toolchain (none in the build environment); never compiled.

//go:build ksareplay
// +build ksareplay

// Package replay evaluates the autoscaler's scaling policy — Run's threshold / cool-down logic
// (main.go:51-77), PodAutoScaler.ScaleUp / ScaleDown (scale/scale.go:54-107) and the attribute sum
// of SqsClient.NumMessages (sqs/sqs.go:56-64) — over many recorded or synthetic queue-depth traces
// at once on a GPU, through the C ABI of libksareplay (include/ksa_replay.h). The live autoscaler
// is untouched; this package sits next to scale/ and sqs/ in the reference module.
//
// Build: the binding is behind the `ksareplay` tag because the reference builds with
// CGO_ENABLED=0 (Dockerfile:4):
//
//	CGO_ENABLED=1 go build -tags ksareplay ./replay ./cmd/ksa-replay
//
// with include/ksa_replay.h in ../include and libksareplay.so in ../lib (or set CGO_CFLAGS /
// CGO_LDFLAGS). Without the tag only reference_test.go is compiled, which needs neither.
package replay

/*
#cgo CFLAGS: -I${SRCDIR}/../include
#cgo LDFLAGS: -L${SRCDIR}/../lib -lksareplay -Wl,-rpath,${SRCDIR}/../lib
#include <stdlib.h>
#include "ksa_replay.h"
*/
import "C"

import (
	"fmt"
	"reflect"
	"runtime"
	"strings"
	"time"
	"unsafe"

	"kube-sqs-autoscaler/scale"
)

// Attribute bits, in the order of sqs.DefaultAttributeNames (sqs/sqs.go:28-32).
const (
	AttrMessages = uint32(C.KSA_ATTR_MESSAGES)
	AttrMessagesDelayed = uint32(C.KSA_ATTR_MESSAGES_DELAYED)
	AttrMessagesNotVisible = uint32(C.KSA_ATTR_MESSAGES_NOT_VISIBLE)
)

// Event bits: one per log line of the reference (include/ksa_replay.h).
const (
	EvUpApplied = uint8(C.KSA_EV_UP_APPLIED)
	EvDownApplied = uint8(C.KSA_EV_DOWN_APPLIED)
	EvUpNoop = uint8(C.KSA_EV_UP_NOOP)
	EvDownNoop = uint8(C.KSA_EV_DOWN_NOOP)
	EvUpCool = uint8(C.KSA_EV_UP_COOL)
	EvDownCool = uint8(C.KSA_EV_DOWN_COOL)
	EvUpFailed = uint8(C.KSA_EV_UP_FAILED)
	EvDownFailed = uint8(C.KSA_EV_DOWN_FAILED)
	EvPollFailed = uint8(C.KSA_EV_POLL_FAILED)
)

var attributeNames = []string{
	"ApproximateNumberOfMessages",
	"ApproximateNumberOfMessagesDelayed",
	"ApproximateNumberOfMessagesNotVisible",
}

// Policy mirrors the flag globals of main.go:16-25 and the fields of scale.PodAutoScaler
// (scale/scale.go:21-29). Pod counts stay Go ints: the library truncates them with int32(...)
// exactly where scale.go does.
type Policy struct {
	PollInterval        time.Duration // --poll-period           main.go:83
	ScaleDownCoolPeriod time.Duration // --scale-down-cool-down  main.go:84
	ScaleUpCoolPeriod   time.Duration // --scale-up-cool-down    main.go:85
	ScaleUpMessages     int           // --scale-up-messages     main.go:86
	ScaleDownMessages   int           // --scale-down-messages   main.go:87
	ScaleUpPods         int           // --scale-up-pods         main.go:88
	ScaleDownPods       int           // --scale-down-pods       main.go:89
	MaxPods             int           // --max-pods              main.go:90
	MinPods             int           // --min-pods              main.go:91
	AttrMask            uint32        // --attribute-names       main.go:93 (Attr* bits)
}

// DefaultPolicy returns the flag defaults of main.go:83-93.
func DefaultPolicy() Policy {
	return Policy{
		PollInterval: 5 * time.Second, ScaleDownCoolPeriod: 30 * time.Second, ScaleUpCoolPeriod: 10 * time.Second,
		ScaleUpMessages: 100, ScaleDownMessages: 10, ScaleUpPods: 1, ScaleDownPods: 1, MaxPods: 5, MinPods: 1,
		AttrMask: AttrMessages | AttrMessagesDelayed | AttrMessagesNotVisible,
	}
}

// FromAutoScaler copies the clamp parameters out of an existing *scale.PodAutoScaler.
func FromAutoScaler(p *scale.PodAutoScaler, base Policy) Policy {
	base.MaxPods, base.MinPods = p.Max, p.Min
	base.ScaleUpPods, base.ScaleDownPods = p.ScaleUpPods, p.ScaleDownPods
	return base
}

// ParseAttributeNames turns the value of --attribute-names (main.go:93,103-110: comma-separated,
// trimmed) into an attribute mask. The planes of a depth array are the selected attributes in the
// order of sqs.DefaultAttributeNames.
func ParseAttributeNames(list string) (uint32, error) {
	var mask uint32
	for _, part := range strings.Split(list, ",") {
		name := strings.TrimSpace(part)
		found := false
		for k, known := range attributeNames {
			if name == known {
				mask |= 1 << uint(k)
				found = true
			}
		}
		if !found {
			return 0, fmt.Errorf("replay: unknown attribute name %q", name)
		}
	}
	return mask, nil
}

// Attrs is the number of attribute planes the policy sums.
func (p Policy) Attrs() int {
	n := 0
	for m := p.AttrMask & 7; m != 0; m &= m - 1 {
		n++
	}
	return n
}

func (p Policy) c() C.ksa_policy {
	return C.ksa_policy{
		scale_up_messages:       C.int64_t(p.ScaleUpMessages),
		scale_down_messages:     C.int64_t(p.ScaleDownMessages),
		max_pods:                C.int64_t(p.MaxPods),
		min_pods:                C.int64_t(p.MinPods),
		scale_up_pods:           C.int64_t(p.ScaleUpPods),
		scale_down_pods:         C.int64_t(p.ScaleDownPods),
		poll_period_ns:          C.int64_t(p.PollInterval),
		scale_up_cool_down_ns:   C.int64_t(p.ScaleUpCoolPeriod),
		scale_down_cool_down_ns: C.int64_t(p.ScaleDownCoolPeriod),
		attr_mask:               C.uint32_t(p.AttrMask),
	}
}

// call runs f on a locked OS thread and turns its status into an error: ksa_last_error() is
// thread-local, so the status and the message must be read on the thread that made the call.
func call(f func() C.int) error {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	rc := f()
	if rc == C.KSA_OK {
		return nil
	}
	return fmt.Errorf("ksareplay: status %d: %s", int(rc), C.GoString(C.ksa_last_error()))
}

// CoolTicks returns the two cool-downs in poll ticks, max(0, ceil(cool/poll)): the form in which
// lastScaleXTime.Add(cool).After(time.Now()) (main.go:52,66) is replayed.
func (p Policy) CoolTicks() (up, down int64, err error) {
	cp := p.c()
	var u, d C.int64_t
	err = call(func() C.int { return C.ksa_policy_cool_ticks(&cp, &u, &d) })
	return int64(u), int64(d), err
}

// Summary is ksa_summary: what one trace's replay did, without the per-tick output.
type Summary struct {
	FinalReplicas, MinReplicas, MaxReplicas                   int32
	UpApplied, DownApplied, UpNoop, DownNoop, UpCool, DownCool uint32
	PollFailed, UpFailed, DownFailed                          uint32
	SumReplicas                                               int64
	Hash                                                      uint64
}

// SummaryBytes views a []Summary as the bytes of the ksa_summary records it holds (64 each), for
// CopyToHost.
func SummaryBytes(s []Summary) []byte {
	if len(s) == 0 {
		return nil
	}
	var b []byte
	h := (*reflect.SliceHeader)(unsafe.Pointer(&b))
	h.Data, h.Len, h.Cap = uintptr(unsafe.Pointer(&s[0])), len(s)*64, len(s)*64
	return b
}

// AllocPinned returns a []byte over page-locked host memory (cudaMallocHost): copies to and from
// it run at full PCIe rate. Free it with FreePinned; the Go garbage collector does not own it.
func AllocPinned(bytes int) ([]byte, error) {
	var p unsafe.Pointer
	if err := call(func() C.int { return C.ksa_host_alloc(&p, C.size_t(bytes)) }); err != nil {
		return nil, err
	}
	var b []byte
	h := (*reflect.SliceHeader)(unsafe.Pointer(&b))
	h.Data, h.Len, h.Cap = uintptr(p), bytes, bytes
	return b, nil
}

// FreePinned releases a slice from AllocPinned.
func FreePinned(b []byte) error {
	if len(b) == 0 {
		return nil
	}
	return call(func() C.int { return C.ksa_host_free(unsafe.Pointer(&b[0])) })
}

// PinSlice page-locks memory the caller already has (ksa_host_register), for instance a slice of a
// shared mapping that the processes of one box, one per GPU, write their shards' summaries into.
// The slice should start and end on page boundaries. Undo with UnpinSlice before unmapping.
func PinSlice(b []byte) error {
	if len(b) == 0 {
		return nil
	}
	return call(func() C.int { return C.ksa_host_register(unsafe.Pointer(&b[0]), C.size_t(len(b))) })
}

// UnpinSlice undoes PinSlice.
func UnpinSlice(b []byte) error {
	if len(b) == 0 {
		return nil
	}
	return call(func() C.int { return C.ksa_host_unregister(unsafe.Pointer(&b[0])) })
}

// Pipeline owns the device staging for up to N traces x A attributes on one GPU
// (ksa_pipeline_create): host slices in, host slices out, copies overlapped with the kernels.
type Pipeline struct {
	h     *C.ksa_pipeline
	n     int
	attrs int
}

// NewPipeline creates a pipeline on `device` moving chunkTicks ticks per stage.
func NewPipeline(device, n, attrs, chunkTicks int) (*Pipeline, error) {
	var h *C.ksa_pipeline
	err := call(func() C.int {
		return C.ksa_pipeline_create(&h, C.int(device), C.int64_t(n), C.int32_t(attrs), C.int64_t(chunkTicks))
	})
	if err != nil {
		return nil, err
	}
	return &Pipeline{h: h, n: n, attrs: attrs}, nil
}

// Close frees the device staging.
func (p *Pipeline) Close() error {
	h := p.h
	p.h = nil
	return call(func() C.int { return C.ksa_pipeline_destroy(h) })
}

// State is the carry state of N traces: Spec.Replicas and the two stamps of main.go:37-38 as
// absolute tick indices (0 = process start). Replay advances it in place.
type State struct {
	Replicas []int32
	LastUp   []int64
	LastDown []int64
}

// NewState is a fresh process for n deployments that all start at `replicas`.
func NewState(n int, replicas int32) State {
	s := State{Replicas: make([]int32, n), LastUp: make([]int64, n), LastDown: make([]int64, n)}
	for i := range s.Replicas {
		s.Replicas[i] = replicas
	}
	return s
}

// Replay runs T ticks for N = len(st.Replicas) traces. depth is [T][A][N] row-major, A =
// pol.Attrs(). replicasOut / eventsOut ([T][N]) may be nil; summaries ([N]) may be nil; fail
// ([T][N] of KSA_FAIL_* bits: 1 poll, 2 ScaleUp, 4 ScaleDown) may be nil for a world without
// errors. tick0 is the number of ticks these traces have already been replayed for.
func (p *Pipeline) Replay(pol Policy, depth []int64, fail []uint8, T int, tick0 int64, st State,
	replicasOut []int32, eventsOut []uint8, summaries []Summary) error {
	n := len(st.Replicas)
	if n == 0 || n > p.n || len(st.LastUp) != n || len(st.LastDown) != n {
		return fmt.Errorf("replay: state of %d traces does not fit a pipeline for %d", n, p.n)
	}
	if pol.Attrs() != p.attrs || len(depth) != T*p.attrs*n {
		return fmt.Errorf("replay: depth has %d values, want T*A*N = %d*%d*%d", len(depth), T, p.attrs, n)
	}
	if (replicasOut != nil && len(replicasOut) != T*n) || (eventsOut != nil && len(eventsOut) != T*n) ||
		(summaries != nil && len(summaries) != n) || (fail != nil && len(fail) != T*n) {
		return fmt.Errorf("replay: an output or the failure mask has the wrong length")
	}
	cp := pol.c()
	var rp *C.int32_t
	var ep *C.uint8_t
	var sp *C.ksa_summary
	if replicasOut != nil {
		rp = (*C.int32_t)(unsafe.Pointer(&replicasOut[0]))
	}
	if eventsOut != nil {
		ep = (*C.uint8_t)(unsafe.Pointer(&eventsOut[0]))
	}
	if summaries != nil {
		sp = (*C.ksa_summary)(unsafe.Pointer(&summaries[0]))
	}
	dp := (*C.int64_t)(unsafe.Pointer(&depth[0]))
	sr := (*C.int32_t)(unsafe.Pointer(&st.Replicas[0]))
	su := (*C.int64_t)(unsafe.Pointer(&st.LastUp[0]))
	sd := (*C.int64_t)(unsafe.Pointer(&st.LastDown[0]))
	return call(func() C.int {
		if fail != nil {
			return C.ksa_pipeline_replay_faults(p.h, &cp, dp, (*C.uint8_t)(unsafe.Pointer(&fail[0])), C.int64_t(T),
				C.int64_t(n), C.int64_t(tick0), sr, su, sd, rp, ep, sp)
		}
		return C.ksa_pipeline_replay(p.h, &cp, dp, C.int64_t(T), C.int64_t(n), C.int64_t(tick0), sr, su, sd, rp, ep, sp)
	})
}

// DeviceCount is the number of CUDA devices the library sees.
func DeviceCount() (int, error) {
	n := int(C.ksa_device_count())
	if n < 0 {
		return 0, fmt.Errorf("ksareplay: status %d: %s", n, C.GoString(C.ksa_last_error()))
	}
	return n, nil
}

// ---- device-resident calls: for hosts that keep the depth arrays on the GPU ------------------------
//
// Device pointers are plain addresses that Go never dereferences; they come from DevAlloc and are
// passed back to the library only.

// DevicePtr is a CUDA device address.
type DevicePtr unsafe.Pointer

// SetDevice selects the GPU for the calling OS thread (pair it with runtime.LockOSThread).
func SetDevice(device int) error { return call(func() C.int { return C.ksa_set_device(C.int(device)) }) }

// DevAlloc allocates device memory (cudaMalloc) on the current device.
func DevAlloc(bytes int) (DevicePtr, error) {
	var p unsafe.Pointer
	err := call(func() C.int { return C.ksa_dev_alloc(&p, C.size_t(bytes)) })
	return DevicePtr(p), err
}

// DevFree releases memory from DevAlloc.
func DevFree(p DevicePtr) error { return call(func() C.int { return C.ksa_dev_free(unsafe.Pointer(p)) }) }

// DeviceState is ksa_state: three device arrays of N elements.
type DeviceState struct{ Replicas, LastUp, LastDown DevicePtr }

func (s DeviceState) c() C.ksa_state {
	return C.ksa_state{replicas_dev: (*C.int32_t)(unsafe.Pointer(s.Replicas)), last_up_tick_dev: (*C.int64_t)(unsafe.Pointer(s.LastUp)),
		last_down_tick_dev: (*C.int64_t)(unsafe.Pointer(s.LastDown))}
}

// GenParams is ksa_gen_params (the synthetic trace generator).
type GenParams struct {
	Seed                                uint64
	Family                              int32 // 0 bands, 1 random walk, 2 adversarial
	LowMax, HighMin, HighMax, WalkStep  int64
	OverflowLog2                        uint32
}

func (g GenParams) c() C.ksa_gen_params {
	return C.ksa_gen_params{seed: C.uint64_t(g.Seed), family: C.int32_t(g.Family), low_max: C.int64_t(g.LowMax),
		high_min: C.int64_t(g.HighMin), high_max: C.int64_t(g.HighMax), walk_step: C.int64_t(g.WalkStep),
		overflow_log2: C.uint32_t(g.OverflowLog2)}
}

// GenTraces writes synthetic depths [T][A][ld] for ticks tick0+1..tick0+T of traces trace0.. into depth.
func GenTraces(g GenParams, tick0 int64, T, A int, trace0 int64, n, ld int, depth, walk DevicePtr) error {
	cg := g.c()
	return call(func() C.int {
		return C.ksa_gen_traces(&cg, C.int64_t(tick0), C.int64_t(T), C.int32_t(A), C.int64_t(trace0), C.int64_t(n),
			C.int64_t(ld), (*C.int64_t)(unsafe.Pointer(depth)), (*C.int64_t)(unsafe.Pointer(walk)), nil)
	})
}

// StateInit fills a fresh carry state: replicas lo..hi by a hash of (seed, trace id), stamps 0.
func StateInit(seed uint64, trace0 int64, n int, lo, hi int32, st DeviceState) error {
	cs := st.c()
	return call(func() C.int {
		return C.ksa_state_init(C.uint64_t(seed), C.int64_t(trace0), C.int64_t(n), C.int32_t(lo), C.int32_t(hi), &cs, nil)
	})
}

// ReplayFull is ksa_replay_full on device-resident arrays (default stream); replicas / events may be nil.
func ReplayFull(pol Policy, depth DevicePtr, T, A, n, ld int, tick0 int64, st DeviceState, replicas, events DevicePtr) error {
	cp, cs := pol.c(), st.c()
	return call(func() C.int {
		return C.ksa_replay_full(&cp, (*C.int64_t)(unsafe.Pointer(depth)), C.int64_t(T), C.int32_t(A), C.int64_t(n), C.int64_t(ld),
			C.int64_t(tick0), &cs, (*C.int32_t)(unsafe.Pointer(replicas)), (*C.uint8_t)(unsafe.Pointer(events)), nil)
	})
}

// ReplaySummary is ksa_replay_summary; SummaryClear starts a fresh accumulation.
func ReplaySummary(pol Policy, depth DevicePtr, T, A, n, ld int, tick0 int64, st DeviceState, summaries DevicePtr) error {
	cp, cs := pol.c(), st.c()
	return call(func() C.int {
		return C.ksa_replay_summary(&cp, (*C.int64_t)(unsafe.Pointer(depth)), C.int64_t(T), C.int32_t(A), C.int64_t(n), C.int64_t(ld),
			C.int64_t(tick0), &cs, (*C.ksa_summary)(unsafe.Pointer(summaries)), nil)
	})
}

// SummaryClear resets n device summaries.
func SummaryClear(summaries DevicePtr, n int) error {
	return call(func() C.int { return C.ksa_summary_clear((*C.ksa_summary)(unsafe.Pointer(summaries)), C.int64_t(n), nil) })
}

// ReplaySweep is ksa_replay_sweep_static: the same traces under up to 4096 policies, one summary
// per (policy, trace) in summaries [P][N]; the kernel-side policy table lives inside the library,
// so a Go host has no device buffer to keep alive for it.
func ReplaySweep(pols []Policy, depth DevicePtr, T, A, n, ld int, replicas0, summaries DevicePtr) error {
	if len(pols) == 0 {
		return fmt.Errorf("replay: empty policy grid")
	}
	cps := make([]C.ksa_policy, len(pols))
	for i, p := range pols {
		cps[i] = p.c()
	}
	return call(func() C.int {
		return C.ksa_replay_sweep_static(&cps[0], C.int32_t(len(cps)), (*C.int64_t)(unsafe.Pointer(depth)), C.int64_t(T), C.int32_t(A),
			C.int64_t(n), C.int64_t(ld), (*C.int32_t)(unsafe.Pointer(replicas0)), (*C.ksa_summary)(unsafe.Pointer(summaries)), nil)
	})
}

// CopyToHost copies device memory into a host slice and waits for it.
func CopyToHost(dst []byte, src DevicePtr) error {
	if len(dst) == 0 {
		return nil
	}
	if err := call(func() C.int {
		return C.ksa_memcpy_d2h(unsafe.Pointer(&dst[0]), unsafe.Pointer(src), C.size_t(len(dst)), nil)
	}); err != nil {
		return err
	}
	return call(func() C.int { return C.ksa_stream_sync(nil) })
}
