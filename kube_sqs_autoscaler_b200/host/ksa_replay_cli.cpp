// ksa-replay — command-line host over the C ABI (include/ksa_replay.h), C++ because the reference
// is compiled code and this image has no Go toolchain. It keeps the reference's operator API: the
// flags of main.go:83-97 with the same names, defaults and Go duration syntax, so a command line
// written for kube-sqs-autoscaler parses unchanged (the flags that only configure the AWS and
// Kubernetes clients are accepted and ignored). Instead of polling one queue forever it replays
// the policy over --traces synthetic queue-depth traces x --ticks polls, sharded over --gpus GPUs
// (one host thread per GPU, no collective, host-side gather of the per-trace summaries), and prints
// what the autoscaler would have done. With --mode sweep every tunable flag of main.go:83-91 may be
// a comma-separated list; the cross product is replayed as a policy grid (ksa_replay_sweep), scored
// (ksa_sweep_reduce, an extension: the reference has no cost model) and the cheapest policies are
// printed as ready-to-paste flag sets.
//
// Links only libksareplay.so: all CUDA calls happen behind the ABI.
#include <algorithm>
#include <cctype>
#include <chrono>
#include <cinttypes>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "ksa_replay.h"

namespace {

[[noreturn]] void die(const std::string& msg) {
  std::fprintf(stderr, "ksa-replay: %s\n", msg.c_str());
  std::exit(2);
}

void check(int rc, const char* what) {
  if (rc != KSA_OK) die(std::string(what) + ": " + ksa_last_error());
}

// time.ParseDuration (what flag.DurationVar uses, main.go:83-85), restated from Go's algorithm:
//   [-+]?([0-9]*(\.[0-9]*)?[a-z]+)+   units ns us µs μs ms s m h; "0" alone is allowed.
// Integer arithmetic in uint64 with Go's overflow checks (leadingInt: a number above 2^63 is an
// error; value * unit and the running total must stay <= 2^63, and only a negative duration may
// be exactly 2^63 ns); the fractional part goes through float64 exactly as in Go:
// uint64(float64(f) * (float64(unit) / scale)), f and scale from leadingFraction, which stops
// taking digits once f would pass 2^63 - 1.
bool parse_duration(const std::string& text, int64_t* out) {
  const uint64_t kLimit = 1ull << 63;
  size_t p = 0;
  const size_t n = text.size();
  bool neg = false;
  if (p < n && (text[p] == '+' || text[p] == '-')) neg = text[p++] == '-';
  if (text.compare(p, std::string::npos, "0") == 0) {
    *out = 0;
    return true;
  }
  if (p >= n) return false;
  uint64_t d = 0;
  while (p < n) {
    if (!(text[p] == '.' || (text[p] >= '0' && text[p] <= '9'))) return false;
    // leadingInt
    uint64_t v = 0;
    const size_t int_begin = p;
    while (p < n && text[p] >= '0' && text[p] <= '9') {
      if (v > kLimit / 10) return false;
      v = v * 10 + (uint64_t)(text[p] - '0');
      if (v > kLimit) return false;
      ++p;
    }
    const bool pre = p != int_begin;
    // leadingFraction
    uint64_t f = 0;
    double scale = 1;
    bool post = false;
    if (p < n && text[p] == '.') {
      ++p;
      const size_t frac_begin = p;
      bool overflow = false;
      while (p < n && text[p] >= '0' && text[p] <= '9') {
        if (!overflow) {
          if (f > (kLimit - 1) / 10) {
            overflow = true;
          } else {
            const uint64_t y = f * 10 + (uint64_t)(text[p] - '0');
            if (y > kLimit) {
              overflow = true;
            } else {
              f = y;
              scale *= 10;
            }
          }
        }
        ++p;
      }
      post = p != frac_begin;
    }
    if (!pre && !post) return false;  // no digits (e.g. ".s")
    const size_t unit_begin = p;
    while (p < n && text[p] != '.' && !(text[p] >= '0' && text[p] <= '9')) ++p;
    if (p == unit_begin) return false;  // missing unit
    const std::string unit = text.substr(unit_begin, p - unit_begin);
    uint64_t mult;
    if (unit == "ns") mult = 1;
    else if (unit == "us" || unit == "\xC2\xB5s" || unit == "\xCE\xBCs") mult = 1000ull;  // U+00B5, U+03BC
    else if (unit == "ms") mult = 1000000ull;
    else if (unit == "s") mult = 1000000000ull;
    else if (unit == "m") mult = 60ull * 1000000000ull;
    else if (unit == "h") mult = 3600ull * 1000000000ull;
    else return false;
    if (v > kLimit / mult) return false;
    v *= mult;
    if (f > 0) {
      v += (uint64_t)((double)f * ((double)mult / scale));
      if (v > kLimit) return false;
    }
    d += v;
    if (d > kLimit) return false;
  }
  if (neg) {
    *out = (int64_t)(0ull - d);  // 2^63 becomes math.MinInt64, as in Go
    return true;
  }
  if (d > kLimit - 1) return false;
  *out = (int64_t)d;
  return true;
}

uint32_t attr_mask(const std::string& names) {
  static const char* kNames[3] = {"ApproximateNumberOfMessages", "ApproximateNumberOfMessagesDelayed",
                                  "ApproximateNumberOfMessagesNotVisible"};  // sqs/sqs.go:28-32
  uint32_t mask = 0;
  size_t p = 0;
  while (p <= names.size()) {
    size_t q = names.find(',', p);
    if (q == std::string::npos) q = names.size();
    std::string n = names.substr(p, q - p);
    n.erase(0, n.find_first_not_of(" \t"));  // strings.TrimSpace, main.go:108
    n.erase(n.find_last_not_of(" \t") + 1);
    bool found = false;
    for (int a = 0; a < 3; ++a)
      if (n == kNames[a]) {
        mask |= 1u << a;
        found = true;
      }
    if (!found) die("unknown attribute name '" + n + "'");
    p = q + 1;
  }
  return mask;
}

struct Options {
  ksa_policy policy;
  int64_t traces = 1 << 20, ticks = 1000;
  uint64_t seed = 0x5EED0001;
  std::string mode = "summary";
  int device = 0;  // first device
  int gpus = 1;    // devices device .. device+gpus-1, one host thread each
  // injected API failures (main.go:44-47,57-60,71-74): each call fails with probability 2^-k, 0 = never
  uint32_t poll_fail_log2 = 0, up_fail_log2 = 0, down_fail_log2 = 0;
  bool faults() const { return poll_fail_log2 || up_fail_log2 || down_fail_log2; }
  // --mode sweep: the value lists of the nine tunables (a single value outside sweep mode) and the cost
  std::vector<int64_t> up_cool, down_cool, up_msgs, down_msgs, up_pods, down_pods, max_pods, min_pods;
  ksa_cost cost = {1, 0, 0, 0};
  int top = 5;
  // recorded depths instead of synthetic ones: a raw file of [A][traces][ticks] values, one queue's
  // history per row (trace-major), as int64 / int32 or as fixed-width NUL-padded decimal strings
  std::string depth_file, depth_format = "i64";
};

Options parse(int argc, char** argv) {
  Options o;
  check(ksa_policy_default(&o.policy), "ksa_policy_default");
  std::map<std::string, std::string> kv;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    if (a.rfind("--", 0) == 0) a = a.substr(2);
    else if (a.rfind("-", 0) == 0) a = a.substr(1);  // Go's flag package accepts -flag too
    else die("unexpected argument '" + a + "'");
    std::string v;
    const size_t eq = a.find('=');
    if (eq != std::string::npos) {
      v = a.substr(eq + 1);
      a = a.substr(0, eq);
    } else if (a == "help" || a == "h") {
      v = "";
    } else {
      if (i + 1 >= argc) die("flag needs an argument: -" + a);
      v = argv[++i];
    }
    kv[a] = v;
  }
  const bool sweep = kv.count("mode") && kv["mode"] == "sweep";
  for (const auto& [k, v] : kv) {
    auto as_int = [&](const char* what) {
      char* end = nullptr;
      const long long x = std::strtoll(v.c_str(), &end, 0);
      if (v.empty() || *end) die(std::string("invalid value \"") + v + "\" for flag -" + what);
      return (int64_t)x;
    };
    auto as_dur = [&](const char* what) {
      int64_t d;
      if (!parse_duration(v, &d)) die(std::string("invalid value \"") + v + "\" for flag -" + what + ": time: invalid duration");
      return d;
    };
    // A tunable: one value, or in sweep mode a comma-separated list of values.
    auto as_list = [&](const char* what, bool duration, std::vector<int64_t>* out) {
      out->clear();
      size_t p = 0;
      while (p <= v.size()) {
        size_t q = sweep ? v.find(',', p) : std::string::npos;
        if (q == std::string::npos) q = v.size();
        const std::string item = v.substr(p, q - p);
        int64_t x;
        if (duration) {
          if (!parse_duration(item, &x)) die(std::string("invalid value \"") + item + "\" for flag -" + what + ": time: invalid duration");
        } else {
          char* end = nullptr;
          x = (int64_t)std::strtoll(item.c_str(), &end, 0);
          if (item.empty() || *end) die(std::string("invalid value \"") + item + "\" for flag -" + what);
        }
        out->push_back(x);
        p = q + 1;
      }
      return out->front();
    };
    if (k == "parse-duration") {  // print what time.ParseDuration gives for the value and stop (no GPU needed)
      std::printf("%lld\n", (long long)as_dur("parse-duration"));
      std::exit(0);
    } else if (k == "poll-period") o.policy.poll_period_ns = as_dur("poll-period");
    else if (k == "scale-down-cool-down") o.policy.scale_down_cool_down_ns = as_list("scale-down-cool-down", true, &o.down_cool);
    else if (k == "scale-up-cool-down") o.policy.scale_up_cool_down_ns = as_list("scale-up-cool-down", true, &o.up_cool);
    else if (k == "scale-up-messages") o.policy.scale_up_messages = as_list("scale-up-messages", false, &o.up_msgs);
    else if (k == "scale-down-messages") o.policy.scale_down_messages = as_list("scale-down-messages", false, &o.down_msgs);
    else if (k == "scale-up-pods") o.policy.scale_up_pods = as_list("scale-up-pods", false, &o.up_pods);
    else if (k == "scale-down-pods") o.policy.scale_down_pods = as_list("scale-down-pods", false, &o.down_pods);
    else if (k == "max-pods") o.policy.max_pods = as_list("max-pods", false, &o.max_pods);
    else if (k == "min-pods") o.policy.min_pods = as_list("min-pods", false, &o.min_pods);
    else if (k == "cost-pod-ticks") o.cost.w_pod_ticks = as_int("cost-pod-ticks");
    else if (k == "cost-saturated") o.cost.w_saturated = as_int("cost-saturated");
    else if (k == "cost-churn") o.cost.w_churn = as_int("cost-churn");
    else if (k == "cost-blocked") o.cost.w_blocked = as_int("cost-blocked");
    else if (k == "top") o.top = (int)as_int("top");
    else if (k == "depth-file") o.depth_file = v;
    else if (k == "depth-format") o.depth_format = v;
    else if (k == "attribute-names") o.policy.attr_mask = attr_mask(v);
    else if (k == "aws-region" || k == "sqs-queue-url" || k == "kubernetes-deployment" || k == "kubernetes-namespace") {
      // network clients of the live daemon (main.go:92,95-97): nothing to do in a replay
    } else if (k == "traces") o.traces = as_int("traces");
    else if (k == "ticks") o.ticks = as_int("ticks");
    else if (k == "seed") o.seed = (uint64_t)as_int("seed");
    else if (k == "mode") o.mode = v;
    else if (k == "device") o.device = (int)as_int("device");
    else if (k == "gpus") o.gpus = (int)as_int("gpus");
    else if (k == "poll-fail-log2") o.poll_fail_log2 = (uint32_t)as_int("poll-fail-log2");
    else if (k == "up-fail-log2") o.up_fail_log2 = (uint32_t)as_int("up-fail-log2");
    else if (k == "down-fail-log2") o.down_fail_log2 = (uint32_t)as_int("down-fail-log2");
    else if (k == "help" || k == "h") {
      std::puts("usage: ksa-replay [reference flags: --poll-period --scale-up-cool-down --scale-down-cool-down\n"
                "  --scale-up-messages --scale-down-messages --scale-up-pods --scale-down-pods --max-pods --min-pods\n"
                "  --attribute-names] [--traces N] [--ticks T] [--seed S] [--mode full|summary|sweep] [--device D] [--gpus G]\n"
                "  --mode sweep: every tunable may be a comma-separated list; [--cost-pod-ticks W] [--cost-saturated W]\n"
                "  [--cost-churn W] [--cost-blocked W] [--top K]\n"
                "  [--depth-file F --depth-format i64|i32|text16|text32]  recorded depths, [attribute][trace][tick], instead of synthetic ones\n"
                "  [--poll-fail-log2 K] [--up-fail-log2 K] [--down-fail-log2 K]  (each API call fails with probability 2^-K)");
      std::exit(0);
    } else {
      die("flag provided but not defined: -" + k);
    }
  }
  if (o.mode != "full" && o.mode != "summary" && o.mode != "sweep") die("--mode must be full, summary or sweep");
  if (o.mode == "sweep" && o.faults()) die("--mode sweep does not take --*-fail-log2");
  if (o.top < 1) die("--top must be >= 1");
  if (o.depth_format != "i64" && o.depth_format != "i32" && o.depth_format != "text16" && o.depth_format != "text32")
    die("--depth-format must be i64, i32, text16 or text32");
  if (!o.depth_file.empty() && o.faults()) die("--depth-file does not take --*-fail-log2");
  if (!o.depth_file.empty() && o.mode == "sweep" && o.depth_format[0] == 't')
    die("--mode sweep reads recordings as --depth-format i64 or i32 (the sweep has no failed polls)");
  if (o.poll_fail_log2 > 63 || o.up_fail_log2 > 63 || o.down_fail_log2 > 63) die("--*-fail-log2 must be in [0, 63]");
  return o;
}

// One GPU's share of the job: traces [begin, end) in N-chunks of at most 2^20 traces (so that a
// chunk's depth array fits next to its outputs whatever --traces is), each chunk in T-chunks with
// the carry state. Summaries land in `host` at their global trace ids; returns seconds spent in
// replay kernels on this device.
double run_shard(const Options& o, int device, int64_t begin, int64_t end, std::vector<ksa_summary>* host) {
  check(ksa_set_device(device), "ksa_set_device");
  const int A = __builtin_popcount(o.policy.attr_mask);
  const int64_t T = o.ticks, NC = std::min<int64_t>(end - begin, 1 << 20), ld = (NC + 15) / 16 * 16;
  const bool full = o.mode == "full";
  const bool faults = o.faults();
  const int64_t chunk = std::min<int64_t>(T, std::max<int64_t>(1, (int64_t)((full ? 48.0e9 : 64.0e9) / ((8.0 * A + (full ? 5 : 0) + (faults ? 1 : 0)) * ld))));
  void* fail = nullptr;
  void *depth = nullptr, *r = nullptr, *ju = nullptr, *jd = nullptr, *summ = nullptr, *rep = nullptr, *evs = nullptr;
  void *r2 = nullptr, *ju2 = nullptr, *jd2 = nullptr;
  check(ksa_dev_alloc(&depth, (size_t)chunk * A * ld * 8), "depth");
  if (faults) check(ksa_dev_alloc(&fail, (size_t)chunk * ld), "fail");
  check(ksa_dev_alloc(&r, (size_t)ld * 4), "state");
  check(ksa_dev_alloc(&ju, (size_t)ld * 8), "state");
  check(ksa_dev_alloc(&jd, (size_t)ld * 8), "state");
  check(ksa_dev_alloc(&summ, (size_t)ld * sizeof(ksa_summary)), "summary");
  ksa_state st = {(int32_t*)r, (int64_t*)ju, (int64_t*)jd};
  ksa_state st_sum = st;
  if (full) {  // per-tick outputs, and a second copy of the carry state for the summary pass
    check(ksa_dev_alloc(&rep, (size_t)chunk * ld * 4), "replicas");
    check(ksa_dev_alloc(&evs, (size_t)chunk * ld), "events");
    check(ksa_dev_alloc(&r2, (size_t)ld * 4), "state");
    check(ksa_dev_alloc(&ju2, (size_t)ld * 8), "state");
    check(ksa_dev_alloc(&jd2, (size_t)ld * 8), "state");
    st_sum = {(int32_t*)r2, (int64_t*)ju2, (int64_t*)jd2};
  }
  const int32_t lo = (int32_t)o.policy.min_pods;
  const int32_t hi = (int32_t)o.policy.max_pods >= lo ? (int32_t)o.policy.max_pods : lo;
  ksa_gen_params g;
  std::memset(&g, 0, sizeof(g));
  g.seed = o.seed;
  g.family = KSA_GEN_BANDS;
  g.low_max = o.policy.scale_down_messages < 0 ? 0 : o.policy.scale_down_messages;
  g.high_min = o.policy.scale_up_messages > g.low_max ? o.policy.scale_up_messages : g.low_max + 1;
  g.high_max = 4 * g.high_min;
  double replay_s = 0;
  for (int64_t c0 = begin; c0 < end; c0 += NC) {
    const int64_t n = std::min(NC, end - c0);
    check(ksa_state_init(o.seed, c0, n, lo, hi, &st, nullptr), "ksa_state_init");
    if (full) check(ksa_state_init(o.seed, c0, n, lo, hi, &st_sum, nullptr), "ksa_state_init");
    check(ksa_summary_clear((ksa_summary*)summ, n, nullptr), "ksa_summary_clear");
    for (int64_t t0 = 0; t0 < T; t0 += chunk) {
      const int64_t tc = std::min(chunk, T - t0);
      check(ksa_gen_traces(&g, t0, tc, A, c0, n, ld, (int64_t*)depth, nullptr, nullptr), "ksa_gen_traces");
      if (faults)
        check(ksa_gen_faults(o.seed, o.poll_fail_log2, o.up_fail_log2, o.down_fail_log2, t0, tc, c0, n, ld, (uint8_t*)fail, nullptr), "ksa_gen_faults");
      check(ksa_stream_sync(nullptr), "sync");
      const int64_t* d = (const int64_t*)depth;
      const uint8_t* f = (const uint8_t*)fail;
      const auto a = std::chrono::steady_clock::now();
      if (full && faults) check(ksa_replay_full_faults(&o.policy, d, f, tc, A, n, ld, t0, &st, (int32_t*)rep, (uint8_t*)evs, nullptr), "ksa_replay_full_faults");
      else if (full) check(ksa_replay_full(&o.policy, d, tc, A, n, ld, t0, &st, (int32_t*)rep, (uint8_t*)evs, nullptr), "ksa_replay_full");
      else if (faults) check(ksa_replay_summary_faults(&o.policy, d, f, tc, A, n, ld, t0, &st, (ksa_summary*)summ, nullptr), "ksa_replay_summary_faults");
      else check(ksa_replay_summary(&o.policy, d, tc, A, n, ld, t0, &st, (ksa_summary*)summ, nullptr), "ksa_replay_summary");
      check(ksa_stream_sync(nullptr), "sync");
      replay_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - a).count();
      if (full && faults) check(ksa_replay_summary_faults(&o.policy, d, f, tc, A, n, ld, t0, &st_sum, (ksa_summary*)summ, nullptr), "ksa_replay_summary_faults");
      else if (full) check(ksa_replay_summary(&o.policy, d, tc, A, n, ld, t0, &st_sum, (ksa_summary*)summ, nullptr), "ksa_replay_summary");
    }
    // host gather: this chunk's summaries at their global trace ids
    check(ksa_memcpy_d2h(host->data() + c0, summ, (size_t)n * sizeof(ksa_summary), nullptr), "d2h");
    check(ksa_stream_sync(nullptr), "sync");
  }
  for (void* p : {depth, fail, r, ju, jd, summ, rep, evs, r2, ju2, jd2})
    if (p) ksa_dev_free(p);
  return replay_s;
}

// ---- --depth-file: recorded queue depths ----------------------------------------------------------

// Reads N-chunks of a recording into the tick-major device layout. The file holds, per selected
// attribute, one row per trace (a queue's history, the order metrics exports come in); rows of
// adjacent traces are adjacent, so an N-chunk of one attribute is one contiguous read. A chunk goes
// host -> device, through ksa_ingest_decimal when the values are strings (strconv.Atoi as
// sqs/sqs.go:58; a string that does not parse fails that poll), and through ksa_ingest_transpose.
struct RecordingLoader {
  const Options& o;
  int A;
  int64_t T, N, elem, max_n;
  bool text;
  FILE* f = nullptr;
  void *hostbuf = nullptr, *src = nullptr, *vals = nullptr, *fsrc = nullptr;

  static int64_t elem_bytes(const std::string& fmt) { return fmt == "i64" ? 8 : fmt == "i32" ? 4 : fmt == "text16" ? 16 : 32; }

  RecordingLoader(const Options& opt, int64_t max_traces)
      : o(opt), A(__builtin_popcount(opt.policy.attr_mask)), T(opt.ticks), N(opt.traces),
        elem(elem_bytes(opt.depth_format)), max_n(max_traces), text(opt.depth_format[0] == 't') {
    f = std::fopen(o.depth_file.c_str(), "rb");
    if (!f) die("cannot open " + o.depth_file);
    check(ksa_host_alloc(&hostbuf, (size_t)max_n * T * elem), "host buffer");
    check(ksa_dev_alloc(&src, (size_t)max_n * T * elem), "source");
    if (text) {
      check(ksa_dev_alloc(&vals, (size_t)max_n * T * 8), "values");
      check(ksa_dev_alloc(&fsrc, (size_t)max_n * T), "parse failures");
    }
  }
  ~RecordingLoader() {
    if (f) std::fclose(f);
    for (void* p : {src, vals, fsrc})
      if (p) ksa_dev_free(p);
    if (hostbuf) ksa_host_free(hostbuf);
  }
  // traces [c0, c0 + n) -> depth [T][A][ld]; fail [T][ld] when the format is text (else untouched)
  void load(int64_t c0, int64_t n, int64_t ld, void* depth, void* fail) {
    for (int a = 0; a < A; ++a) {
      const long long off = ((long long)a * N + c0) * T * elem;
      if (fseeko(f, (off_t)off, SEEK_SET) != 0 || std::fread(hostbuf, (size_t)elem, (size_t)(n * T), f) != (size_t)(n * T))
        die(o.depth_file + ": short read (needs " + std::to_string((long long)A * N * T * elem) + " bytes)");
      check(ksa_memcpy_h2d(src, hostbuf, (size_t)n * T * elem, nullptr), "h2d");
      check(ksa_stream_sync(nullptr), "sync");
      if (text) {
        check(ksa_ingest_decimal((const char*)src, (int32_t)elem, n * T, (int64_t*)vals, (uint8_t*)fsrc, nullptr), "ksa_ingest_decimal");
        check(ksa_ingest_transpose(vals, 8, T, T, n, A, a, ld, (int64_t*)depth, nullptr), "ksa_ingest_transpose");
        check(ksa_ingest_transpose_mask((const uint8_t*)fsrc, T, T, n, ld, (uint8_t*)fail, a > 0, nullptr), "ksa_ingest_transpose_mask");
      } else {
        check(ksa_ingest_transpose(src, (int32_t)elem, T, T, n, A, a, ld, (int64_t*)depth, nullptr), "ksa_ingest_transpose");
      }
    }
    check(ksa_stream_sync(nullptr), "sync");
  }
};

// One GPU's share of a recorded block: load a chunk, then the same replay calls as the synthetic path.
double run_file_shard(const Options& o, int device, int64_t begin, int64_t end, std::vector<ksa_summary>* host) {
  check(ksa_set_device(device), "ksa_set_device");
  const int A = __builtin_popcount(o.policy.attr_mask);
  const int64_t T = o.ticks;
  const bool text = o.depth_format[0] == 't';
  const int64_t elem = RecordingLoader::elem_bytes(o.depth_format);
  const bool full = o.mode == "full";
  const double per_trace = (double)T * (elem + (text ? 9 : 0) + 8.0 * A + (text ? 1 : 0) + (full ? 5 : 0)) + 256;
  int64_t NC = std::max<int64_t>(16, (int64_t)(48.0e9 / per_trace) / 16 * 16);
  NC = std::min<int64_t>(std::min<int64_t>(NC, 1 << 20), (end - begin + 15) / 16 * 16);
  const int64_t ld = NC;
  RecordingLoader loader(o, NC);
  void *depth = nullptr, *fail = nullptr;
  void *r = nullptr, *ju = nullptr, *jd = nullptr, *r2 = nullptr, *ju2 = nullptr, *jd2 = nullptr, *summ = nullptr;
  void *rep = nullptr, *evs = nullptr;
  if (text) check(ksa_dev_alloc(&fail, (size_t)T * ld), "fail mask");
  check(ksa_dev_alloc(&depth, (size_t)T * A * ld * 8), "depth");
  check(ksa_dev_alloc(&r, (size_t)ld * 4), "state");
  check(ksa_dev_alloc(&ju, (size_t)ld * 8), "state");
  check(ksa_dev_alloc(&jd, (size_t)ld * 8), "state");
  check(ksa_dev_alloc(&summ, (size_t)ld * sizeof(ksa_summary)), "summary");
  ksa_state st = {(int32_t*)r, (int64_t*)ju, (int64_t*)jd};
  ksa_state st_sum = st;
  if (full) {
    check(ksa_dev_alloc(&rep, (size_t)T * ld * 4), "replicas");
    check(ksa_dev_alloc(&evs, (size_t)T * ld), "events");
    check(ksa_dev_alloc(&r2, (size_t)ld * 4), "state");
    check(ksa_dev_alloc(&ju2, (size_t)ld * 8), "state");
    check(ksa_dev_alloc(&jd2, (size_t)ld * 8), "state");
    st_sum = {(int32_t*)r2, (int64_t*)ju2, (int64_t*)jd2};
  }
  const int32_t lo = (int32_t)o.policy.min_pods;
  const int32_t hi = (int32_t)o.policy.max_pods >= lo ? (int32_t)o.policy.max_pods : lo;
  double replay_s = 0;
  for (int64_t c0 = begin; c0 < end; c0 += NC) {
    const int64_t n = std::min(NC, end - c0);
    loader.load(c0, n, ld, depth, fail);
    check(ksa_state_init(o.seed, c0, n, lo, hi, &st, nullptr), "ksa_state_init");
    if (full) check(ksa_state_init(o.seed, c0, n, lo, hi, &st_sum, nullptr), "ksa_state_init");
    check(ksa_summary_clear((ksa_summary*)summ, n, nullptr), "ksa_summary_clear");
    check(ksa_stream_sync(nullptr), "sync");
    const int64_t* d = (const int64_t*)depth;
    const uint8_t* fm = (const uint8_t*)fail;
    const auto t_a = std::chrono::steady_clock::now();
    if (full && text) check(ksa_replay_full_faults(&o.policy, d, fm, T, A, n, ld, 0, &st, (int32_t*)rep, (uint8_t*)evs, nullptr), "ksa_replay_full_faults");
    else if (full) check(ksa_replay_full(&o.policy, d, T, A, n, ld, 0, &st, (int32_t*)rep, (uint8_t*)evs, nullptr), "ksa_replay_full");
    else if (text) check(ksa_replay_summary_faults(&o.policy, d, fm, T, A, n, ld, 0, &st, (ksa_summary*)summ, nullptr), "ksa_replay_summary_faults");
    else check(ksa_replay_summary(&o.policy, d, T, A, n, ld, 0, &st, (ksa_summary*)summ, nullptr), "ksa_replay_summary");
    check(ksa_stream_sync(nullptr), "sync");
    replay_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t_a).count();
    if (full && text) check(ksa_replay_summary_faults(&o.policy, d, fm, T, A, n, ld, 0, &st_sum, (ksa_summary*)summ, nullptr), "ksa_replay_summary_faults");
    else if (full) check(ksa_replay_summary(&o.policy, d, T, A, n, ld, 0, &st_sum, (ksa_summary*)summ, nullptr), "ksa_replay_summary");
    check(ksa_memcpy_d2h(host->data() + c0, summ, (size_t)n * sizeof(ksa_summary), nullptr), "d2h");
    check(ksa_stream_sync(nullptr), "sync");
  }
  for (void* p : {depth, fail, r, ju, jd, r2, ju2, jd2, summ, rep, evs})
    if (p) ksa_dev_free(p);
  return replay_s;
}

// ---- --mode sweep -------------------------------------------------------------------------------

// The policy grid: cross product of the value lists, the last tunable varying fastest.
std::vector<ksa_policy> build_grid(const Options& o) {
  auto one = [](const std::vector<int64_t>& v, int64_t dflt) { return v.empty() ? std::vector<int64_t>{dflt} : v; };
  const auto ucs = one(o.up_cool, o.policy.scale_up_cool_down_ns), dcs = one(o.down_cool, o.policy.scale_down_cool_down_ns);
  const auto ums = one(o.up_msgs, o.policy.scale_up_messages), dms = one(o.down_msgs, o.policy.scale_down_messages);
  const auto ups = one(o.up_pods, o.policy.scale_up_pods), dps = one(o.down_pods, o.policy.scale_down_pods);
  const auto mxs = one(o.max_pods, o.policy.max_pods), mns = one(o.min_pods, o.policy.min_pods);
  std::vector<ksa_policy> grid;
  for (int64_t um : ums) for (int64_t dm : dms) for (int64_t uc : ucs) for (int64_t dc : dcs)
  for (int64_t up : ups) for (int64_t dp : dps) for (int64_t mx : mxs) for (int64_t mn : mns) {
    ksa_policy p = o.policy;
    p.scale_up_messages = um; p.scale_down_messages = dm;
    p.scale_up_cool_down_ns = uc; p.scale_down_cool_down_ns = dc;
    p.scale_up_pods = up; p.scale_down_pods = dp; p.max_pods = mx; p.min_pods = mn;
    grid.push_back(p);
    if (grid.size() > (1u << 20)) die("policy grid larger than 2^20");
  }
  return grid;
}

// One GPU's share of a sweep: traces [begin, end) in N-chunks sized so that the P x n summaries fit
// in 8 GB; per chunk: generate, sweep, score. policy_cost accumulates over the chunks on the device
// (ksa_sweep_reduce adds); best_policy lands in `best` at the global trace ids.
double run_sweep_shard(const Options& o, const std::vector<ksa_policy>& grid, int device, int64_t begin, int64_t end,
                       std::vector<int64_t>* cost_out, std::vector<int32_t>* best) {
  check(ksa_set_device(device), "ksa_set_device");
  const int A = __builtin_popcount(o.policy.attr_mask);
  const int32_t P = (int32_t)grid.size();
  const int64_t T = o.ticks;
  int64_t NC = std::min<int64_t>(end - begin, std::max<int64_t>(16, (int64_t)(8.0e9 / (64.0 * P)) / 16 * 16));
  NC = std::min<int64_t>(NC, std::max<int64_t>(16, (int64_t)(32.0e9 / (8.0 * A * T)) / 16 * 16));
  const int64_t ld = (NC + 15) / 16 * 16;
  void *depth = nullptr, *r = nullptr, *ju = nullptr, *jd = nullptr, *summ = nullptr, *table = nullptr;
  void *pcost = nullptr, *bpol = nullptr, *bcost = nullptr;
  check(ksa_dev_alloc(&depth, (size_t)T * A * ld * 8), "depth");
  check(ksa_dev_alloc(&r, (size_t)ld * 4), "state");
  check(ksa_dev_alloc(&ju, (size_t)ld * 8), "state");
  check(ksa_dev_alloc(&jd, (size_t)ld * 8), "state");
  check(ksa_dev_alloc(&summ, (size_t)P * NC * sizeof(ksa_summary)), "summary");
  check(ksa_dev_alloc(&table, (size_t)P * 48), "policy table");
  check(ksa_dev_alloc(&pcost, (size_t)P * 8), "policy cost");
  check(ksa_dev_alloc(&bpol, (size_t)NC * 4), "best policy");
  check(ksa_dev_alloc(&bcost, (size_t)NC * 8), "best cost");
  std::vector<int64_t> zeros((size_t)P, 0);
  check(ksa_memcpy_h2d(pcost, zeros.data(), (size_t)P * 8, nullptr), "h2d");
  ksa_state st = {(int32_t*)r, (int64_t*)ju, (int64_t*)jd};
  const int32_t lo = (int32_t)o.policy.min_pods;
  const int32_t hi = (int32_t)o.policy.max_pods >= lo ? (int32_t)o.policy.max_pods : lo;
  ksa_gen_params g;
  std::memset(&g, 0, sizeof(g));
  g.seed = o.seed;
  g.family = KSA_GEN_WALK;  // queue depths that drift: the shape a tuner cares about
  g.low_max = 0;            // (the band limits are unused by the walk but validated)
  g.high_min = 1;
  g.high_max = 4 * std::max<int64_t>(1, o.policy.scale_up_messages);
  g.walk_step = std::max<int64_t>(1, g.high_max / 32);
  void* walk = nullptr;
  check(ksa_dev_alloc(&walk, (size_t)ld * 8), "walk state");
  const bool recorded = !o.depth_file.empty();
  RecordingLoader* loader = recorded ? new RecordingLoader(o, NC) : nullptr;
  double secs = 0;
  {  // untimed short pass: the first launch of a kernel pays for loading it onto the device, and
     // an idle GPU needs some tens of milliseconds of work before its clocks are up
    const int64_t n = std::min<int64_t>(NC, end - begin);
    const int64_t tw = std::min<int64_t>(T, 512);
    check(ksa_state_init(o.seed, begin, n, lo, hi, &st, nullptr), "ksa_state_init");
    check(ksa_gen_walk_init(&g, begin, n, (int64_t*)walk, nullptr), "ksa_gen_walk_init");
    check(ksa_gen_traces(&g, 0, tw, A, begin, n, ld, (int64_t*)depth, (int64_t*)walk, nullptr), "ksa_gen_traces");
    check(ksa_replay_sweep(grid.data(), P, table, (const int64_t*)depth, tw, A, n, ld, (const int32_t*)r,
                           (ksa_summary*)summ, nullptr), "ksa_replay_sweep");
    check(ksa_sweep_reduce((const ksa_summary*)summ, P, n, &o.cost, (int64_t*)pcost, (int32_t*)bpol, (int64_t*)bcost, nullptr),
          "ksa_sweep_reduce");
    check(ksa_memcpy_h2d(pcost, zeros.data(), (size_t)P * 8, nullptr), "h2d");
    check(ksa_stream_sync(nullptr), "sync");
  }
  for (int64_t c0 = begin; c0 < end; c0 += NC) {
    const int64_t n = std::min(NC, end - c0);
    check(ksa_state_init(o.seed, c0, n, lo, hi, &st, nullptr), "ksa_state_init");
    if (recorded) {
      loader->load(c0, n, ld, depth, nullptr);
    } else {
      check(ksa_gen_walk_init(&g, c0, n, (int64_t*)walk, nullptr), "ksa_gen_walk_init");
      check(ksa_gen_traces(&g, 0, T, A, c0, n, ld, (int64_t*)depth, (int64_t*)walk, nullptr), "ksa_gen_traces");
    }
    check(ksa_stream_sync(nullptr), "sync");
    const auto a = std::chrono::steady_clock::now();
    check(ksa_replay_sweep(grid.data(), P, table, (const int64_t*)depth, T, A, n, ld, (const int32_t*)r,
                           (ksa_summary*)summ, nullptr), "ksa_replay_sweep");
    check(ksa_sweep_reduce((const ksa_summary*)summ, P, n, &o.cost, (int64_t*)pcost, (int32_t*)bpol, (int64_t*)bcost, nullptr),
          "ksa_sweep_reduce");
    check(ksa_stream_sync(nullptr), "sync");
    secs += std::chrono::duration<double>(std::chrono::steady_clock::now() - a).count();
    check(ksa_memcpy_d2h(best->data() + c0, bpol, (size_t)n * 4, nullptr), "d2h");
    check(ksa_stream_sync(nullptr), "sync");
  }
  check(ksa_memcpy_d2h(cost_out->data(), pcost, (size_t)P * 8, nullptr), "d2h");
  check(ksa_stream_sync(nullptr), "sync");
  for (void* p : {depth, r, ju, jd, summ, table, pcost, bpol, bcost, walk}) ksa_dev_free(p);
  delete loader;
  return secs;
}

std::string fmt_duration(int64_t ns) {
  char buf[64];
  if (ns % 1000000000 == 0) std::snprintf(buf, sizeof buf, "%" PRId64 "s", ns / 1000000000);
  else if (ns % 1000000 == 0) std::snprintf(buf, sizeof buf, "%" PRId64 "ms", ns / 1000000);
  else std::snprintf(buf, sizeof buf, "%" PRId64 "ns", ns);
  return buf;
}

int run_sweep(const Options& o) {
  const std::vector<ksa_policy> grid = build_grid(o);
  const int32_t P = (int32_t)grid.size();
  const int64_t N = o.traces, T = o.ticks;
  const int have = ksa_device_count();
  if (have < 0) die(std::string("ksa_device_count: ") + ksa_last_error());
  if (o.gpus < 1 || o.device + o.gpus > have)
    die("--gpus " + std::to_string(o.gpus) + " from --device " + std::to_string(o.device) + ": only " + std::to_string(have) + " CUDA devices");
  // Shards of traces, one host thread per GPU; the per-policy totals of the shards are added on the host.
  std::vector<std::vector<int64_t>> costs((size_t)o.gpus, std::vector<int64_t>((size_t)P, 0));
  std::vector<int32_t> best((size_t)N, -1);
  std::vector<double> secs((size_t)o.gpus, 0.0);
  std::vector<std::thread> workers;
  for (int g = 0; g < o.gpus; ++g) {
    const int64_t base = N / o.gpus, extra = N % o.gpus;
    const int64_t begin = g * base + std::min<int64_t>(g, extra);
    const int64_t end = begin + base + (g < extra ? 1 : 0);
    workers.emplace_back([&, g, begin, end] {
      if (end > begin) secs[(size_t)g] = run_sweep_shard(o, grid, o.device + g, begin, end, &costs[(size_t)g], &best);
    });
  }
  for (auto& w : workers) w.join();
  std::vector<int64_t> total((size_t)P, 0);
  for (const auto& c : costs)
    for (int32_t p = 0; p < P; ++p) total[(size_t)p] = (int64_t)((uint64_t)total[(size_t)p] + (uint64_t)c[(size_t)p]);
  std::vector<int64_t> wins((size_t)P, 0);
  for (int32_t b : best) if (b >= 0) ++wins[(size_t)b];
  std::vector<int32_t> order((size_t)P);
  for (int32_t p = 0; p < P; ++p) order[(size_t)p] = p;
  std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return total[(size_t)a] < total[(size_t)b]; });
  const double s = *std::max_element(secs.begin(), secs.end());
  std::printf("swept %d policies x %" PRId64 " traces x %" PRId64 " ticks on %d GPU(s) in %.3f ms (slowest GPU, sweep + scoring): %.3e decisions/s\n",
              P, N, T, o.gpus, s * 1e3, (double)P * (double)N * (double)T / s);
  std::printf("cost = %" PRId64 "*pod_ticks + %" PRId64 "*ticks_saturated_at_max + %" PRId64 "*scale_events + %" PRId64 "*cool_down_skips\n",
              o.cost.w_pod_ticks, o.cost.w_saturated, o.cost.w_churn, o.cost.w_blocked);
  for (int k = 0; k < std::min<int>(o.top, P); ++k) {
    const ksa_policy& p = grid[(size_t)order[(size_t)k]];
    std::printf("#%d policy %d cost %" PRId64 " best-for %" PRId64 " traces: --scale-up-messages=%" PRId64 " --scale-down-messages=%" PRId64
                " --scale-up-cool-down=%s --scale-down-cool-down=%s --scale-up-pods=%" PRId64 " --scale-down-pods=%" PRId64
                " --max-pods=%" PRId64 " --min-pods=%" PRId64 "\n",
                k + 1, order[(size_t)k], total[(size_t)order[(size_t)k]], wins[(size_t)order[(size_t)k]], p.scale_up_messages,
                p.scale_down_messages, fmt_duration(p.scale_up_cool_down_ns).c_str(), fmt_duration(p.scale_down_cool_down_ns).c_str(),
                p.scale_up_pods, p.scale_down_pods, p.max_pods, p.min_pods);
  }
  return 0;
}

}  // namespace

int main(int argc, char** argv) {
  Options o = parse(argc, argv);
  if (o.mode == "sweep") return run_sweep(o);
  const int A = __builtin_popcount(o.policy.attr_mask);
  const int64_t N = o.traces, T = o.ticks;
  int64_t cu = 0, cd = 0;
  check(ksa_policy_cool_ticks(&o.policy, &cu, &cd), "ksa_policy_cool_ticks");
  const int have = ksa_device_count();
  if (have < 0) die(std::string("ksa_device_count: ") + ksa_last_error());
  if (o.gpus < 1 || o.device + o.gpus > have)
    die("--gpus " + std::to_string(o.gpus) + " from --device " + std::to_string(o.device) + ": only " + std::to_string(have) + " CUDA devices");

  // Traces are independent: contiguous shards, one host thread per GPU, no collective; the only
  // cross-GPU step is the host-side gather of the per-trace summaries.
  std::vector<ksa_summary> host((size_t)N);
  std::vector<double> secs((size_t)o.gpus, 0.0);
  std::vector<std::thread> workers;
  for (int g = 0; g < o.gpus; ++g) {
    const int64_t base = N / o.gpus, extra = N % o.gpus;
    const int64_t begin = g * base + std::min<int64_t>(g, extra);
    const int64_t end = begin + base + (g < extra ? 1 : 0);
    workers.emplace_back([&, g, begin, end] {
      if (end > begin)
        secs[(size_t)g] = o.depth_file.empty() ? run_shard(o, o.device + g, begin, end, &host)
                                               : run_file_shard(o, o.device + g, begin, end, &host);
    });
  }
  for (auto& w : workers) w.join();
  const double replay_s = *std::max_element(secs.begin(), secs.end());

  uint64_t tot[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, digest = 0;
  long double replica_ticks = 0;
  for (const ksa_summary& s : host) {
    tot[0] += s.up_applied; tot[1] += s.down_applied; tot[2] += s.up_noop;
    tot[3] += s.down_noop; tot[4] += s.up_cool; tot[5] += s.down_cool;
    tot[6] += s.poll_failed; tot[7] += s.up_failed; tot[8] += s.down_failed;
    replica_ticks += (long double)s.sum_replicas;
    digest ^= s.hash;
  }
  std::printf("policy: up>=%" PRId64 " down<=%" PRId64 " steps +%" PRId64 "/-%" PRId64 " pods [%" PRId64 ",%" PRId64
              "] cool-down ticks up %" PRId64 " down %" PRId64 " attributes %d\n",
              o.policy.scale_up_messages, o.policy.scale_down_messages, o.policy.scale_up_pods, o.policy.scale_down_pods,
              o.policy.min_pods, o.policy.max_pods, cu, cd, A);
  std::printf("replayed %" PRId64 " traces x %" PRId64 " ticks (%s mode) on %d GPU(s) in %.3f ms of kernel wall time (slowest GPU): %.3e decisions/s\n",
              N, T, o.mode.c_str(), o.gpus, replay_s * 1e3, (double)N * (double)T / replay_s);
  std::printf("scale up successful %" PRIu64 " | scale down successful %" PRIu64 " | at max %" PRIu64 " | at min %" PRIu64
              " | up cool-down skips %" PRIu64 " | down cool-down skips %" PRIu64 "\n",
              tot[0], tot[1], tot[2], tot[3], tot[4], tot[5]);
  if (o.faults() || (!o.depth_file.empty() && o.depth_format[0] == 't'))
    std::printf("failed to get SQS messages %" PRIu64 " | failed scaling up %" PRIu64 " | failed scaling down %" PRIu64 "\n",
                tot[6], tot[7], tot[8]);
  std::printf("mean replicas %.4Lf | xor of trajectory fingerprints %016" PRIx64 "\n",
              replica_ticks / ((long double)N * (long double)T), digest);
  return 0;
}
