// UNVERIFIED: written without a Go toolchain (none in the build environment); never compiled.

//go:build ksareplay
// +build ksareplay

// ksa-replay replays the autoscaler's policy over synthetic queue-depth traces on a GPU. It takes
// the reference's own flags (main.go:83-97) with their names, defaults and units, so the flag set
// of a live deployment can be pasted here unchanged; --aws-region, --sqs-queue-url,
// --kubernetes-deployment and --kubernetes-namespace are accepted and ignored (there is no queue
// and no cluster in a replay).
package main

import (
	"flag"
	"fmt"
	"os"
	"runtime"
	"time"

	"kube-sqs-autoscaler/replay"
)

var (
	// the reference's globals, main.go:16-33
	pollInterval          time.Duration
	scaleDownCoolPeriod   time.Duration
	scaleUpCoolPeriod     time.Duration
	scaleUpMessages       int
	scaleDownMessages     int
	scaleUpPods           int
	scaleDownPods         int
	maxPods               int
	minPods               int
	awsRegion             string
	attributeNames        string
	defaultAttributeNames string = "ApproximateNumberOfMessages,ApproximateNumberOfMessagesDelayed,ApproximateNumberOfMessagesNotVisible"

	sqsQueueUrl              string
	kubernetesDeploymentName string
	kubernetesNamespace      string

	// replay-only flags
	traces  int
	ticks   int
	seed    uint64
	device  int
	initMin int
	initMax int
)

func die(err error) {
	fmt.Fprintln(os.Stderr, "ksa-replay:", err)
	os.Exit(1)
}

func main() {
	// main.go:83-97, verbatim names / defaults / help texts
	flag.DurationVar(&pollInterval, "poll-period", 5*time.Second, "The interval in seconds for checking if scaling is required")
	flag.DurationVar(&scaleDownCoolPeriod, "scale-down-cool-down", 30*time.Second, "The cool down period for scaling down")
	flag.DurationVar(&scaleUpCoolPeriod, "scale-up-cool-down", 10*time.Second, "The cool down period for scaling up")
	flag.IntVar(&scaleUpMessages, "scale-up-messages", 100, "Number of sqs messages queued up required for scaling up")
	flag.IntVar(&scaleDownMessages, "scale-down-messages", 10, "Number of messages required to scaling down")
	flag.IntVar(&scaleUpPods, "scale-up-pods", 1, "Number of Pod in scaling up")
	flag.IntVar(&scaleDownPods, "scale-down-pods", 1, "Number of Pod in scaling down")
	flag.IntVar(&maxPods, "max-pods", 5, "Max pods that kube-sqs-autoscaler can scale")
	flag.IntVar(&minPods, "min-pods", 1, "Min pods that kube-sqs-autoscaler can scale")
	flag.StringVar(&awsRegion, "aws-region", "", "Your AWS region (ignored by the replay)")
	flag.StringVar(&attributeNames, "attribute-names", defaultAttributeNames, "A comma-separated list of queue attribute names to query in calculating the number of messages")
	flag.StringVar(&sqsQueueUrl, "sqs-queue-url", "", "The sqs queue url (ignored by the replay)")
	flag.StringVar(&kubernetesDeploymentName, "kubernetes-deployment", "", "Kubernetes Deployment to scale (ignored by the replay)")
	flag.StringVar(&kubernetesNamespace, "kubernetes-namespace", "default", "The namespace your deployment is running in (ignored by the replay)")

	flag.IntVar(&traces, "traces", 1<<20, "Number of independent deployments to replay")
	flag.IntVar(&ticks, "ticks", 10000, "Number of polls to replay for each of them")
	flag.Uint64Var(&seed, "seed", 0x5EED0001, "Seed of the synthetic queue depths")
	flag.IntVar(&device, "device", 0, "CUDA device")
	flag.IntVar(&initMin, "init-min", 1, "Smallest initial replica count")
	flag.IntVar(&initMax, "init-max", 5, "Largest initial replica count")
	flag.Parse()

	if pollInterval <= 0 {
		die(fmt.Errorf("--poll-period must be > 0 (Run sleeps one poll period per tick, main.go:41)"))
	}
	mask, err := replay.ParseAttributeNames(attributeNames)
	if err != nil {
		die(err)
	}
	pol := replay.Policy{
		PollInterval: pollInterval, ScaleDownCoolPeriod: scaleDownCoolPeriod, ScaleUpCoolPeriod: scaleUpCoolPeriod,
		ScaleUpMessages: scaleUpMessages, ScaleDownMessages: scaleDownMessages, ScaleUpPods: scaleUpPods,
		ScaleDownPods: scaleDownPods, MaxPods: maxPods, MinPods: minPods, AttrMask: mask,
	}
	attrs := pol.Attrs()

	runtime.LockOSThread() // the CUDA device is a property of the OS thread
	defer runtime.UnlockOSThread()
	if err := replay.SetDevice(device); err != nil {
		die(err)
	}

	n, T := traces, ticks
	alloc := func(bytes int) replay.DevicePtr {
		p, err := replay.DevAlloc(bytes)
		if err != nil {
			die(err)
		}
		return p
	}
	// depth [T][A][N] int64 in chunks of at most chunkTicks ticks: the carry state makes chunking exact
	chunkTicks := T
	for chunkTicks*attrs*n*8 > 64<<30 && chunkTicks > 1 {
		chunkTicks = (chunkTicks + 1) / 2
	}
	depth := alloc(chunkTicks * attrs * n * 8)
	st := replay.DeviceState{Replicas: alloc(n * 4), LastUp: alloc(n * 8), LastDown: alloc(n * 8)}
	summaries := alloc(n * 64)
	if err := replay.StateInit(seed, 0, n, int32(initMin), int32(initMax), st); err != nil {
		die(err)
	}
	if err := replay.SummaryClear(summaries, n); err != nil {
		die(err)
	}
	// bands around the thresholds: 40 % at or under --scale-down-messages, 30 % between, 30 % at or over --scale-up-messages
	gen := replay.GenParams{Seed: seed, Family: 0, LowMax: int64(scaleDownMessages), HighMin: int64(scaleUpMessages),
		HighMax: 4 * int64(scaleUpMessages)}

	began := time.Now()
	for t0 := 0; t0 < T; t0 += chunkTicks {
		tc := chunkTicks
		if T-t0 < tc {
			tc = T - t0
		}
		if err := replay.GenTraces(gen, int64(t0), tc, attrs, 0, n, n, depth, nil); err != nil {
			die(err)
		}
		if err := replay.ReplaySummary(pol, depth, tc, attrs, n, n, int64(t0), st, summaries); err != nil {
			die(err)
		}
	}
	host := make([]replay.Summary, n)
	if err := replay.CopyToHost(replay.SummaryBytes(host), summaries); err != nil { // also waits for the kernels
		die(err)
	}
	elapsed := time.Since(began)

	var tot replay.Summary
	var sumFinal int64
	for i := range host {
		s := &host[i]
		tot.UpApplied += s.UpApplied
		tot.DownApplied += s.DownApplied
		tot.UpNoop += s.UpNoop
		tot.DownNoop += s.DownNoop
		tot.UpCool += s.UpCool
		tot.DownCool += s.DownCool
		tot.SumReplicas += s.SumReplicas
		sumFinal += int64(s.FinalReplicas)
	}
	up, down, _ := pol.CoolTicks()
	fmt.Printf("replayed %d deployments x %d polls (%d attribute(s)); cool-downs %d / %d ticks\n", n, T, attrs, up, down)
	fmt.Printf("  scale up   : %d applied, %d at max (no-op), %d skipped for cool-down\n", tot.UpApplied, tot.UpNoop, tot.UpCool)
	fmt.Printf("  scale down : %d applied, %d at min (no-op), %d skipped for cool-down\n", tot.DownApplied, tot.DownNoop, tot.DownCool)
	fmt.Printf("  mean replicas over the run %.3f, at the end %.3f\n",
		float64(tot.SumReplicas)/float64(n)/float64(T), float64(sumFinal)/float64(n))
	fmt.Printf("  %.3f s including generation: %.3e decisions/s\n", elapsed.Seconds(), float64(n)*float64(T)/elapsed.Seconds())
}
