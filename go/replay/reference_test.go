// UNVERIFIED: written without a Go toolchain (none in the build environment); never compiled.
//
// reference_test.go checks the UNMODIFIED reference — scale.PodAutoScaler (scale/scale.go:54-107)
// on client-go's fake clientset and kubesqs.SqsClient (sqs/sqs.go:45-67) on a mock SQS client, the
// same doubles main_test.go:241-304 uses — against tests/golden/replay_fixtures.json, the cases the
// CUDA kernels and the C / Python restatements of this repository are tested with. It needs no GPU
// and no cgo: it is the step that turns "three restatements of the Go text agree" into "the Go
// code itself agrees".
//
// Run cannot be called as it is: it never returns (main.go:40), sleeps for real (main.go:41) and
// reads the wall clock (main.go:37,38,52,62,66,76). runVirtual below is a copy of its loop body,
// statement by statement, with `now` injected: tick k happens at start + k*pollInterval. ScaleUp,
// ScaleDown and NumMessages are the reference's own, unmodified.
//
// Where to put it: copy this directory into the reference module as kube-sqs-autoscaler/replay
// (next to scale/ and sqs/), then
//
//	KSA_FIXTURES=/path/to/replay_fixtures.json go test ./replay -run TestReferenceMatchesFixtures
//
// (the modules of go.mod:5-19 must be available; nothing else is needed).
package replay

import (
	"context"
	"encoding/json"
	"errors"
	"io/ioutil"
	"os"
	"strconv"
	"strings"
	"testing"
	"time"

	"kube-sqs-autoscaler/scale"
	kubesqs "kube-sqs-autoscaler/sqs"

	"github.com/aws/aws-sdk-go/aws"
	"github.com/aws/aws-sdk-go/service/sqs"
	log "github.com/sirupsen/logrus"
	appsv1 "k8s.io/api/apps/v1"
	metav1 "k8s.io/apimachinery/pkg/apis/meta/v1"
	"k8s.io/apimachinery/pkg/runtime"
	"k8s.io/client-go/kubernetes/fake"
	k8stesting "k8s.io/client-go/testing"
)

// Event bits of include/ksa_replay.h (KSA_EV_*): one per log line of the reference.
const (
	evUpApplied   = 0x01 // "Scale up successful"                        scale/scale.go:77
	evDownApplied = 0x02 // "Scale down successful"                      scale/scale.go:105
	evUpNoop      = 0x04 // "More than max pods running"                 scale/scale.go:63
	evDownNoop    = 0x08 // "Less than min pods running"                 scale/scale.go:90
	evUpCool      = 0x10 // "Waiting for cool down, skipping scale up"   main.go:53
	evDownCool    = 0x20 // "Waiting for cool down, skipping scale down" main.go:67
	evUpFailed    = 0x40 // "Failed scaling up"                          main.go:58
	evDownFailed  = 0x80 // "Failed scaling down"                        main.go:72
	evPollFailed  = 0xC0 // "Failed to get SQS messages"                 main.go:45
)

// Failure bits of a fixture's "fail" entries (KSA_FAIL_*).
const (
	failPoll = 1 // GetQueueAttributes returns an error   sqs/sqs.go:51-54
	failUp   = 2 // the Get inside ScaleUp returns one    scale/scale.go:55-58
	failDown = 4 // the Get inside ScaleDown returns one  scale/scale.go:82-85
)

type fixturePolicy struct {
	PollPeriodNs        int64    `json:"poll_period_ns"`
	ScaleUpCoolDownNs   int64    `json:"scale_up_cool_down_ns"`
	ScaleDownCoolDownNs int64    `json:"scale_down_cool_down_ns"`
	ScaleUpMessages     int      `json:"scale_up_messages"`
	ScaleDownMessages   int      `json:"scale_down_messages"`
	ScaleUpPods         int      `json:"scale_up_pods"`
	ScaleDownPods       int      `json:"scale_down_pods"`
	MaxPods             int      `json:"max_pods"`
	MinPods             int      `json:"min_pods"`
	AttributeNames      []string `json:"attribute_names"`
}

type fixtureCase struct {
	ID        string        `json:"id"`
	Source    string        `json:"source"`
	Policy    fixturePolicy `json:"policy"`
	Replicas0 int32         `json:"replicas0"`
	Depth     [][]int64     `json:"depth"`
	Fail      []uint8       `json:"fail"`
	Replicas  []int32       `json:"replicas"`
	Events    []uint8       `json:"events"`
}

type fixtureFile struct {
	Cases []fixtureCase `json:"cases"`
}

func loadFixtures(t *testing.T) fixtureFile {
	path := os.Getenv("KSA_FIXTURES")
	if path == "" {
		path = "../../tests/golden/replay_fixtures.json" // when run from go/replay inside this repository
	}
	raw, err := ioutil.ReadFile(path) // (os.ReadFile needs Go 1.16; go.mod says 1.15)
	if err != nil {
		t.Skipf("fixtures not found (%v): set KSA_FIXTURES", err)
	}
	var f fixtureFile
	if err := json.Unmarshal(raw, &f); err != nil {
		t.Fatalf("%s: %v", path, err)
	}
	return f
}

// mockSQS is main_test.go:273-286 plus one switch: failNext makes the next GetQueueAttributes
// return an error, which NumMessages wraps and returns (sqs/sqs.go:51-54).
type mockSQS struct {
	QueueAttributes *sqs.GetQueueAttributesOutput
	failNext        bool
}

func (m *mockSQS) GetQueueAttributes(*sqs.GetQueueAttributesInput) (*sqs.GetQueueAttributesOutput, error) {
	if m.failNext {
		m.failNext = false
		return nil, errors.New("injected: GetQueueAttributes failed")
	}
	return m.QueueAttributes, nil
}

func (m *mockSQS) SetQueueAttributes(input *sqs.SetQueueAttributesInput) (*sqs.SetQueueAttributesOutput, error) {
	m.QueueAttributes = &sqs.GetQueueAttributesOutput{Attributes: input.Attributes}
	return &sqs.SetQueueAttributesOutput{}, nil
}

// eventHook turns the reference's own log lines into event bits, so that the comparison covers
// what the autoscaler printed and not only where the replica count ended up.
type eventHook struct{ bits uint8 }

func (h *eventHook) Levels() []log.Level { return log.AllLevels }

func (h *eventHook) Fire(e *log.Entry) error {
	m := e.Message
	switch {
	case strings.HasPrefix(m, "Scale up successful"):
		h.bits |= evUpApplied
	case strings.HasPrefix(m, "Scale down successful"):
		h.bits |= evDownApplied
	case strings.HasPrefix(m, "More than max pods running"):
		h.bits |= evUpNoop
	case strings.HasPrefix(m, "Less than min pods running"):
		h.bits |= evDownNoop
	case strings.HasPrefix(m, "Waiting for cool down, skipping scale up"):
		h.bits |= evUpCool
	case strings.HasPrefix(m, "Waiting for cool down, skipping scale down"):
		h.bits |= evDownCool
	case strings.HasPrefix(m, "Failed scaling up"):
		h.bits |= evUpFailed
	case strings.HasPrefix(m, "Failed scaling down"):
		h.bits |= evDownFailed
	case strings.HasPrefix(m, "Failed to get SQS messages"):
		h.bits |= evPollFailed
	}
	return nil
}

// world is one deployment with its autoscaler and queue: the doubles of main_test.go:241-304, with
// handles kept on the clientset (for the injected Get failure and ClearActions) and the mock.
type world struct {
	clientset *fake.Clientset
	scaler    *scale.PodAutoScaler
	queue     *mockSQS
	sqsClient *kubesqs.SqsClient
	failGet   bool
}

func newWorld(c fixtureCase) *world {
	initialReplicas := c.Replicas0
	w := &world{}
	w.clientset = fake.NewSimpleClientset(&appsv1.Deployment{
		ObjectMeta: metav1.ObjectMeta{Name: "deploy", Namespace: "namespace", Annotations: map[string]string{}},
		Spec:       appsv1.DeploymentSpec{Replicas: &initialReplicas},
	})
	// scale.go's Get comes first in ScaleUp / ScaleDown (:55, :82): an error there is what a
	// KSA_FAIL_UP / KSA_FAIL_DOWN bit stands for.
	w.clientset.PrependReactor("get", "deployments", func(k8stesting.Action) (bool, runtime.Object, error) {
		if w.failGet {
			return true, nil, errors.New("injected: Get failed")
		}
		return false, nil, nil
	})
	w.scaler = &scale.PodAutoScaler{
		Client:        w.clientset.AppsV1().Deployments("namespace"),
		Min:           c.Policy.MinPods,
		Max:           c.Policy.MaxPods,
		ScaleDownPods: c.Policy.ScaleDownPods,
		ScaleUpPods:   c.Policy.ScaleUpPods,
		Deployment:    "deploy",
		Namespace:     "namespace",
	}
	w.queue = &mockSQS{}
	names := make([]*string, 0, len(c.Policy.AttributeNames))
	for _, n := range c.Policy.AttributeNames {
		names = append(names, aws.String(n))
	}
	w.sqsClient = &kubesqs.SqsClient{Client: w.queue, QueueUrl: "example.com", AttributeNames: names}
	return w
}

// runVirtual is the body of Run's loop (main.go:40-78) with the clock injected; the line numbers on
// the right are main.go's. Returns Spec.Replicas and the event bits after every tick.
func runVirtual(t *testing.T, c fixtureCase) ([]int32, []uint8) {
	w := newWorld(c)
	hook := &eventHook{}
	log.AddHook(hook)
	log.SetOutput(nopWriter{})
	defer log.StandardLogger().ReplaceHooks(make(log.LevelHooks))

	pollInterval := time.Duration(c.Policy.PollPeriodNs)
	scaleUpCoolPeriod := time.Duration(c.Policy.ScaleUpCoolDownNs)
	scaleDownCoolPeriod := time.Duration(c.Policy.ScaleDownCoolDownNs)
	scaleUpMessages := c.Policy.ScaleUpMessages
	scaleDownMessages := c.Policy.ScaleDownMessages

	ctx := context.Background()                  // :36
	start := time.Unix(1_600_000_000, 0)         // "time.Now()" at process start
	lastScaleUpTime := start                     // :37
	lastScaleDownTime := start                   // :38
	replicas := make([]int32, 0, len(c.Depth))
	events := make([]uint8, 0, len(c.Depth))

	tick := func(now time.Time, fail uint8) {
		numMessages, err := w.sqsClient.NumMessages() // :43
		if err != nil {                               // :44
			log.Errorf("Failed to get SQS messages: %v", err)
			return // continue, :46
		}
		log.Infof("Found %d messages in the queue", numMessages) // :49

		if numMessages >= scaleUpMessages { // :51
			if lastScaleUpTime.Add(scaleUpCoolPeriod).After(now) { // :52
				log.Info("Waiting for cool down, skipping scale up ")
				return // continue, :54
			}
			w.failGet = fail&failUp != 0
			err := w.scaler.ScaleUp(ctx) // :57
			w.failGet = false
			if err != nil {
				log.Errorf("Failed scaling up: %v", err)
				return // continue, :59
			}
			lastScaleUpTime = now // :62
		}

		if numMessages <= scaleDownMessages { // :65
			if lastScaleDownTime.Add(scaleDownCoolPeriod).After(now) { // :66
				log.Info("Waiting for cool down, skipping scale down")
				return // continue, :68
			}
			w.failGet = fail&failDown != 0
			err := w.scaler.ScaleDown(ctx) // :71
			w.failGet = false
			if err != nil {
				log.Errorf("Failed scaling down: %v", err)
				return // continue, :73
			}
			lastScaleDownTime = now // :76
		}
	}

	for k := 1; k <= len(c.Depth); k++ {
		now := start.Add(time.Duration(k) * pollInterval) // time.Sleep(pollInterval), :41
		attrs := map[string]*string{}
		for a, name := range c.Policy.AttributeNames {
			attrs[name] = aws.String(strconv.FormatInt(c.Depth[k-1][a], 10))
		}
		if _, err := w.queue.SetQueueAttributes(&sqs.SetQueueAttributesInput{Attributes: attrs}); err != nil { // main_test.go:49
			t.Fatal(err)
		}
		var fail uint8
		if c.Fail != nil {
			fail = c.Fail[k-1]
		}
		w.queue.failNext = fail&failPoll != 0
		hook.bits = 0
		tick(now, fail)
		w.queue.failNext = false
		d, err := w.scaler.Client.Get(ctx, "deploy", metav1.GetOptions{})
		if err != nil {
			t.Fatal(err)
		}
		replicas = append(replicas, *d.Spec.Replicas)
		events = append(events, hook.bits)
		if k%100000 == 0 {
			w.clientset.ClearActions() // the fake records every Get / Update
		}
	}
	return replicas, events
}

type nopWriter struct{}

func (nopWriter) Write(p []byte) (int, error) { return len(p), nil }

func TestReferenceMatchesFixtures(t *testing.T) {
	for _, c := range loadFixtures(t).Cases {
		c := c
		t.Run(c.ID, func(t *testing.T) {
			replicas, events := runVirtual(t, c)
			for k := range c.Depth {
				if replicas[k] != c.Replicas[k] || events[k] != c.Events[k] {
					t.Fatalf("%s (%s): tick %d: reference has replicas %d, events %#x; fixture has %d, %#x",
						c.ID, c.Source, k+1, replicas[k], events[k], c.Replicas[k], c.Events[k])
				}
			}
		})
	}
}

// The direct-call tests of scale/scale_test.go do not go through Run; as a sanity check of the
// harness itself, the same sequences through runVirtual's world give the reference's own asserted
// values (scale_test.go:23,27,32: 3 -> 4 -> 5 -> 5).
func TestHarnessReproducesScaleUpTest(t *testing.T) {
	c := fixtureCase{Replicas0: 3, Policy: fixturePolicy{MaxPods: 5, MinPods: 1, ScaleUpPods: 1, ScaleDownPods: 1,
		AttributeNames: []string{"ApproximateNumberOfMessages"}}}
	w := newWorld(c)
	want := []int32{4, 5, 5}
	for i, exp := range want {
		if err := w.scaler.ScaleUp(context.Background()); err != nil {
			t.Fatal(err)
		}
		d, _ := w.scaler.Client.Get(context.Background(), "deploy", metav1.GetOptions{})
		if *d.Spec.Replicas != exp {
			t.Fatalf("call %d: %d, want %d", i+1, *d.Spec.Replicas, exp)
		}
	}
}
