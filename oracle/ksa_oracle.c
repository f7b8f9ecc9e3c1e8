/*
 * ksa_oracle.c — CPU restatement of the reference policy. TEST INFRASTRUCTURE (see ksa_oracle.h
 * for who may call it and for the parity status). Plain C11 + pthreads; written in the reference's
 * own control-flow shape (early returns where Run `continue`s, one function per Go function), NOT
 * in the branch-free shape of the CUDA kernels, so that the two do not share mistakes.
 */
#define _GNU_SOURCE
#include "ksa_oracle.h"
#include "../include/ksa_tracegen.h"

#include <limits.h>
#include <pthread.h>
#include <sched.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

/* ksa_summary.hash (include/ksa_replay.h): two 32-bit lanes, event bits spread to 5-bit fields. */
#define HASH_MUL_LO 0x9E3779B1u
#define HASH_MUL_HI 0x85EBCA6Bu
static uint32_t wide_event(uint32_t ev) {
  uint32_t w = 0;
  for (int k = 0; k < 6; ++k) w |= ((ev >> k) & 1u) << (5 * k);
  return w | (((ev >> 6) & 1u) << 30) | (((ev >> 7) & 1u) << 31);
}
static uint64_t hash_step(uint64_t h, int32_t r, uint32_t ev) {
  const uint32_t lo = (uint32_t)h * HASH_MUL_LO + (uint32_t)r;
  const uint32_t hi = (uint32_t)(h >> 32) * HASH_MUL_HI + wide_event(ev);
  return ((uint64_t)hi << 32) | lo;
}

/* Go's int32(x) on a Go int: keep the low 32 bits (scale/scale.go:62,66-68,89,94-96). */
static inline int32_t go_int32(int64_t x) { return (int32_t)(uint32_t)(uint64_t)x; }
/* Go integer arithmetic wraps; C signed overflow is undefined, so go through unsigned. */
static inline int32_t wrap_add32(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
static inline int32_t wrap_sub32(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }
static inline int64_t wrap_add64(int64_t a, int64_t b) { return (int64_t)((uint64_t)a + (uint64_t)b); }

/* sqs/sqs.go:56-64: `messages = messages + approximateNumberOfMessages` over the selected names. */
int64_t ksa_oracle_num_messages(const int64_t* attrs, int32_t A) {
  int64_t messages = 0; /* sqs.go:56 `var messages int` */
  for (int32_t a = 0; a < A; ++a) messages = wrap_add64(messages, attrs[a]); /* sqs.go:57-64 */
  return messages;
}

/* scale/scale.go:54-79 */
uint8_t ksa_oracle_scale_up(const ksa_policy* p, int32_t* replicas) {
  int32_t current = *replicas;                 /* :55-60 Get + Spec.Replicas          */
  if (current >= go_int32(p->max_pods))        /* :62                                  */
    return KSA_EV_UP_NOOP;                     /* :63-64 log + return nil              */
  int32_t next = wrap_add32(current, go_int32(p->scale_up_pods)); /* :66               */
  if (next > go_int32(p->max_pods))            /* :67                                  */
    next = go_int32(p->max_pods);              /* :68                                  */
  *replicas = next;                            /* :70-72 Update                        */
  return KSA_EV_UP_APPLIED;                    /* :77-78                               */
}

/* scale/scale.go:81-107 */
uint8_t ksa_oracle_scale_down(const ksa_policy* p, int32_t* replicas) {
  int32_t current = *replicas;                 /* :82-87                               */
  if (current <= go_int32(p->min_pods))        /* :89                                  */
    return KSA_EV_DOWN_NOOP;                   /* :90-91                               */
  int32_t next = wrap_sub32(current, go_int32(p->scale_down_pods)); /* :94             */
  if (next < go_int32(p->min_pods))            /* :95                                  */
    next = go_int32(p->min_pods);              /* :96                                  */
  *replicas = next;                            /* :98-100                              */
  return KSA_EV_DOWN_APPLIED;                  /* :105-106                             */
}

/* main.go:35-80, virtual wall clock, one trace. */
void ksa_oracle_run_wallclock(const ksa_policy* p, const int64_t* depth, const uint8_t* fail,
                              int64_t T, int32_t A, int64_t tick0, int32_t* replicas_state,
                              int64_t* last_up_tick, int64_t* last_down_tick, int32_t* replicas_out,
                              uint8_t* events_out) {
  typedef __int128 ns_t;
  const ns_t poll = p->poll_period_ns;
  ns_t last_up = (ns_t)*last_up_tick * poll;     /* main.go:37 (tick 0 = process start) */
  ns_t last_down = (ns_t)*last_down_tick * poll; /* main.go:38                          */
  int32_t r = *replicas_state;
  for (int64_t i = 0; i < T; ++i) {
    const int64_t k = tick0 + 1 + i;
    const ns_t now = (ns_t)k * poll;             /* main.go:41 time.Sleep(pollInterval) */
    const uint8_t f = fail ? fail[i] : 0;
    uint8_t ev = 0;
    int skip = 0;
    if (f & KSA_FAIL_POLL) {                     /* main.go:44 err != nil                */
      ev = KSA_EV_POLL_FAILED;                   /* :45 log                              */
      skip = 2;                                  /* :46 continue                         */
    }
    const int64_t n = ksa_oracle_num_messages(depth + i * A, A); /* main.go:43          */
    if (!skip && n >= p->scale_up_messages) {    /* main.go:51                           */
      if (last_up + (ns_t)p->scale_up_cool_down_ns > now) { /* :52 Add(cool).After(now)  */
        ev |= KSA_EV_UP_COOL;                    /* :53                                  */
        skip = 1;                                /* :54 continue                         */
      } else if (f & KSA_FAIL_UP) {              /* :57 err != nil                       */
        ev |= KSA_EV_UP_FAILED;                  /* :58 log                              */
        skip = 1;                                /* :59 continue (no stamp)              */
      } else {
        ev |= ksa_oracle_scale_up(p, &r);        /* :57                                  */
        last_up = now;                           /* :62 (also after the nil no-op)       */
        *last_up_tick = k;
      }
    }
    if (!skip && n <= p->scale_down_messages) {  /* main.go:65                           */
      if (last_down + (ns_t)p->scale_down_cool_down_ns > now) { /* :66                   */
        ev |= KSA_EV_DOWN_COOL;                  /* :67-68                               */
      } else if (f & KSA_FAIL_DOWN) {            /* :71 err != nil                       */
        ev |= KSA_EV_DOWN_FAILED;                /* :72-73 log, continue (no stamp)      */
      } else {
        ev |= ksa_oracle_scale_down(p, &r);      /* :71                                  */
        last_down = now;                         /* :76                                  */
        *last_down_tick = k;
      }
    }
    if (replicas_out) replicas_out[i] = r;
    if (events_out) events_out[i] = ev;
  }
  *replicas_state = r;
}

int ksa_oracle_cool_ticks(const ksa_policy* p, int64_t* up_ticks, int64_t* down_ticks) {
  if (p->poll_period_ns <= 0) return -1;
  const int64_t poll = p->poll_period_ns;
  const int64_t cu = p->scale_up_cool_down_ns, cd = p->scale_down_cool_down_ns;
  *up_ticks = cu <= 0 ? 0 : cu / poll + (cu % poll != 0);
  *down_ticks = cd <= 0 ? 0 : cd / poll + (cd % poll != 0);
  return 0;
}

/* One tick of main.go:40-78 in the integer-tick model. */
static inline uint8_t tick_model(const ksa_policy* p, int64_t cu, int64_t cd, int64_t n, uint8_t f,
                                 int64_t k, int32_t* r, int64_t* ju, int64_t* jd) {
  uint8_t ev = 0;
  if (f & KSA_FAIL_POLL) return KSA_EV_POLL_FAILED; /* main.go:44-47 */
  if (n >= p->scale_up_messages) {               /* main.go:51    */
    if (k - *ju < cu) return KSA_EV_UP_COOL;     /* main.go:52-55 */
    if (f & KSA_FAIL_UP) return KSA_EV_UP_FAILED; /* main.go:57-60 */
    ev |= ksa_oracle_scale_up(p, r);             /* main.go:57    */
    *ju = k;                                     /* main.go:62    */
  }
  if (n <= p->scale_down_messages) {             /* main.go:65    */
    if (k - *jd < cd) return ev | KSA_EV_DOWN_COOL; /* main.go:66-69 */
    if (f & KSA_FAIL_DOWN) return ev | KSA_EV_DOWN_FAILED; /* main.go:71-74 */
    ev |= ksa_oracle_scale_down(p, r);           /* main.go:71    */
    *jd = k;                                     /* main.go:76    */
  }
  return ev;
}

static inline void summary_add(ksa_summary* s, int32_t r, uint8_t ev) {
  s->final_replicas = r;
  if (r < s->min_replicas) s->min_replicas = r;
  if (r > s->max_replicas) s->max_replicas = r;
  s->up_applied += (ev & KSA_EV_UP_APPLIED) != 0;
  s->down_applied += (ev & KSA_EV_DOWN_APPLIED) != 0;
  s->up_noop += (ev & KSA_EV_UP_NOOP) != 0;
  s->down_noop += (ev & KSA_EV_DOWN_NOOP) != 0;
  s->up_cool += (ev & KSA_EV_UP_COOL) != 0;
  s->down_cool += (ev & KSA_EV_DOWN_COOL) != 0;
  if (ev == KSA_EV_POLL_FAILED) {
    s->poll_failed += 1;
  } else {
    s->up_failed += (ev & KSA_EV_UP_FAILED) != 0;
    s->down_failed += (ev & KSA_EV_DOWN_FAILED) != 0;
  }
  s->sum_replicas = wrap_add64(s->sum_replicas, (int64_t)r);
  s->hash = hash_step(s->hash, r, ev);
}

void ksa_oracle_summary_clear(ksa_summary* s, int64_t N) {
  for (int64_t n = 0; n < N; ++n) {
    memset(&s[n], 0, sizeof(s[n]));
    s[n].min_replicas = INT32_MAX;
    s[n].max_replicas = INT32_MIN;
  }
}

typedef struct {
  const ksa_policy* p;
  const int64_t* depth;
  const uint8_t* fail; /* [T][ld] KSA_FAIL_* bits, or NULL */
  int64_t T, N, ld, tick0, cu, cd;
  int32_t A;
  int32_t* st_r;
  int64_t* st_ju;
  int64_t* st_jd;
  int32_t* replicas;
  uint8_t* events;
  ksa_summary* summary;
  int64_t n0, n1;
} replay_job;

#define BLOCK 2048 /* traces whose state stays in L1 while a worker walks down the ticks */

static void* replay_worker(void* arg) {
  const replay_job* j = (const replay_job*)arg;
  for (int64_t b0 = j->n0; b0 < j->n1; b0 += BLOCK) {
    const int64_t b1 = b0 + BLOCK < j->n1 ? b0 + BLOCK : j->n1;
    for (int64_t t = 0; t < j->T; ++t) {
      const int64_t k = j->tick0 + 1 + t;
      const int64_t* row = j->depth + t * j->A * j->ld;
      for (int64_t n = b0; n < b1; ++n) {
        int64_t attrs[KSA_MAX_ATTRS];
        for (int32_t a = 0; a < j->A; ++a) attrs[a] = row[a * j->ld + n];
        const int64_t msgs = ksa_oracle_num_messages(attrs, j->A);
        const uint8_t f = j->fail ? j->fail[t * j->ld + n] : 0;
        const uint8_t ev =
            tick_model(j->p, j->cu, j->cd, msgs, f, k, &j->st_r[n], &j->st_ju[n], &j->st_jd[n]);
        if (j->replicas) j->replicas[t * j->ld + n] = j->st_r[n];
        if (j->events) j->events[t * j->ld + n] = ev;
        if (j->summary) summary_add(&j->summary[n], j->st_r[n], ev);
      }
    }
  }
  return NULL;
}

int ksa_oracle_online_cpus(void) {
  cpu_set_t set;
  if (sched_getaffinity(0, sizeof(set), &set) == 0) {
    int c = CPU_COUNT(&set);
    if (c > 0) return c;
  }
  long c = sysconf(_SC_NPROCESSORS_ONLN);
  return c > 0 ? (int)c : 1;
}

typedef void* (*worker_fn)(void*);

/* Split [0, N) over `threads` pthreads in BLOCK-aligned ranges; jobs[i] is a copy of *proto with
 * n0/n1 set through the two offsets given. */
static int run_split(worker_fn fn, const void* proto, size_t job_size, size_t off_n0, size_t off_n1,
                     int64_t N, int threads) {
  if (threads <= 0) threads = ksa_oracle_online_cpus();
  int64_t blocks = (N + BLOCK - 1) / BLOCK;
  if (threads > blocks) threads = (int)blocks;
  if (threads < 1) threads = 1;
  char* jobs = (char*)malloc(job_size * (size_t)threads);
  pthread_t* tids = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
  if (!jobs || !tids) {
    free(jobs);
    free(tids);
    return -1;
  }
  for (int i = 0; i < threads; ++i) {
    char* job = jobs + job_size * (size_t)i;
    memcpy(job, proto, job_size);
    int64_t lo = blocks * i / threads * BLOCK, hi = blocks * (i + 1) / threads * BLOCK;
    if (hi > N) hi = N;
    memcpy(job + off_n0, &lo, sizeof(lo));
    memcpy(job + off_n1, &hi, sizeof(hi));
  }
  int started = 0;
  for (int i = 1; i < threads; ++i) {
    if (pthread_create(&tids[i], NULL, fn, jobs + job_size * (size_t)i) != 0) break;
    started = i;
  }
  fn(jobs); /* the calling thread takes range 0 */
  for (int i = started + 1; i < threads; ++i) fn(jobs + job_size * (size_t)i); /* create failed */
  for (int i = 1; i <= started; ++i) pthread_join(tids[i], NULL);
  free(jobs);
  free(tids);
  return threads;
}

int ksa_oracle_replay(const ksa_policy* p, const int64_t* depth, const uint8_t* fail, int64_t T,
                      int32_t A, int64_t N, int64_t ld, int64_t tick0, int32_t* state_replicas, int64_t* state_last_up,
                      int64_t* state_last_down, int32_t* replicas_out, uint8_t* events_out,
                      ksa_summary* summary, int threads) {
  replay_job proto;
  memset(&proto, 0, sizeof(proto));
  if (!p || !depth || T < 0 || N < 0 || ld < N || A < 1 || A > KSA_MAX_ATTRS) return -1;
  if (ksa_oracle_cool_ticks(p, &proto.cu, &proto.cd) != 0) return -1;
  proto.p = p;
  proto.depth = depth;
  proto.fail = fail;
  proto.T = T;
  proto.N = N;
  proto.ld = ld;
  proto.tick0 = tick0;
  proto.A = A;
  proto.st_r = state_replicas;
  proto.st_ju = state_last_up;
  proto.st_jd = state_last_down;
  proto.replicas = replicas_out;
  proto.events = events_out;
  proto.summary = summary;
  if (N == 0 || T == 0) return 1;
  return run_split(replay_worker, &proto, sizeof(proto), offsetof(replay_job, n0),
                   offsetof(replay_job, n1), N, threads);
}

void ksa_oracle_traj_hash(const int32_t* replicas, const uint8_t* events, int64_t T, int64_t N,
                          int64_t ld, int32_t chain, uint64_t* hash) {
  for (int64_t n = 0; n < N; ++n) {
    uint64_t h = chain ? hash[n] : 0;
    for (int64_t t = 0; t < T; ++t)
      h = hash_step(h, replicas[t * ld + n], events[t * ld + n]);
    hash[n] = h;
  }
}

/* ---- generator, host side -------------------------------------------------------------- */

typedef struct {
  const ksa_gen_params* g;
  int64_t tick0, T, trace0, N, ld;
  int32_t A;
  int64_t* depth;
  int64_t* walk;
  int64_t n0, n1;
} gen_job;

static inline void gen_store(const ksa_gen_params* g, int32_t A, int64_t v, uint64_t h,
                             int64_t* cell, int64_t ld) {
  if (A == 1) {
    cell[0] = v;
    return;
  }
  int64_t parts[3];
  ksa_gen_split3(g, v, h, parts);
  if (A == 2) {
    cell[0] = parts[0];
    cell[ld] = wrap_add64(parts[1], parts[2]);
  } else {
    cell[0] = parts[0];
    cell[ld] = parts[1];
    cell[2 * ld] = parts[2];
  }
}

static void* gen_worker(void* arg) {
  const gen_job* j = (const gen_job*)arg;
  const ksa_gen_params* g = j->g;
  for (int64_t b0 = j->n0; b0 < j->n1; b0 += BLOCK) {
    const int64_t b1 = b0 + BLOCK < j->n1 ? b0 + BLOCK : j->n1;
    uint64_t keys[BLOCK];
    for (int64_t n = b0; n < b1; ++n)
      keys[n - b0] = ksa_trace_key(g->seed, 0, (uint64_t)(j->trace0 + n));
    for (int64_t t = 0; t < j->T; ++t) {
      const uint64_t k = (uint64_t)(j->tick0 + 1 + t);
      int64_t* row = j->depth + t * j->A * j->ld;
      for (int64_t n = b0; n < b1; ++n) {
        const uint64_t h = ksa_entry_word(keys[n - b0], k);
        int64_t v;
        if (g->family == KSA_GEN_WALK) {
          v = ksa_gen_walk_step(g, j->walk[n], h);
          j->walk[n] = v;
        } else {
          v = ksa_gen_iid_value(g, h);
        }
        gen_store(g, j->A, v, h, row + n, j->ld);
      }
    }
  }
  return NULL;
}

int ksa_oracle_gen_traces(const ksa_gen_params* g, int64_t tick0, int64_t T, int32_t A,
                          int64_t trace0, int64_t N, int64_t ld, int64_t* depth, int64_t* walk,
                          int threads) {
  if (!g || !depth || T < 0 || N < 0 || ld < N || A < 1 || A > KSA_MAX_ATTRS) return -1;
  if (g->family == KSA_GEN_WALK && !walk) return -1;
  gen_job proto;
  memset(&proto, 0, sizeof(proto));
  proto.g = g;
  proto.tick0 = tick0;
  proto.T = T;
  proto.trace0 = trace0;
  proto.N = N;
  proto.ld = ld;
  proto.A = A;
  proto.depth = depth;
  proto.walk = walk;
  if (N == 0 || T == 0) return 1;
  return run_split(gen_worker, &proto, sizeof(proto), offsetof(gen_job, n0), offsetof(gen_job, n1),
                   N, threads);
}

void ksa_oracle_gen_walk_init(const ksa_gen_params* g, int64_t trace0, int64_t N, int64_t* walk) {
  for (int64_t n = 0; n < N; ++n)
    walk[n] = ksa_gen_walk_start(g, ksa_trace_key(g->seed, 0, (uint64_t)(trace0 + n)));
}

void ksa_oracle_state_init(uint64_t seed, int64_t trace0, int64_t N, int32_t lo, int32_t hi,
                           int32_t* replicas, int64_t* last_up, int64_t* last_down) {
  for (int64_t n = 0; n < N; ++n) {
    replicas[n] = ksa_gen_replicas0(seed, (uint64_t)(trace0 + n), lo, hi);
    last_up[n] = 0;
    last_down[n] = 0;
  }
}

void ksa_oracle_gen_faults(uint64_t seed, uint32_t poll_log2, uint32_t up_log2, uint32_t down_log2,
                           int64_t tick0, int64_t T, int64_t trace0, int64_t N, int64_t ld,
                           uint8_t* fail) {
  for (int64_t n = 0; n < N; ++n) {
    const uint64_t key = ksa_fault_key(seed, (uint64_t)(trace0 + n));
    for (int64_t t = 0; t < T; ++t)
      fail[t * ld + n] = ksa_gen_fault_byte(key, (uint64_t)(tick0 + 1 + t), poll_log2, up_log2, down_log2);
  }
}

/* ---- strconv.Atoi, as sqs/sqs.go:58 applies it to every attribute value --------------------
 * Restated from the published Go 1.15 strconv algorithm (the stdlib is not part of the reference
 * tree): Atoi(s) = ParseInt(s, 10, 0): optional sign, then ParseUint: every byte must be '0'..'9'
 * (no underscores in base 10), at least one digit, overflow of uint64 or of the signed range is
 * ErrRange, anything else ErrSyntax. Both error kinds make NumMessages return an error
 * (sqs/sqs.go:59-61). Records are fixed-width, NUL padded. Returns the number of failed records. */
int64_t ksa_oracle_atoi(const char* text, int32_t width, int64_t count, int64_t* values, uint8_t* fail) {
  int64_t bad = 0;
  for (int64_t i = 0; i < count; ++i) {
    const unsigned char* s = (const unsigned char*)text + i * width;
    int len = 0;
    while (len < width && s[len] != 0) ++len;
    int ok = 1;
    int neg = 0;
    int p = 0;
    if (len == 0) ok = 0;                                  /* ParseInt: s == "" -> syntaxError */
    if (ok && (s[0] == '+' || s[0] == '-')) { neg = s[0] == '-'; p = 1; }
    if (ok && p == len) ok = 0;                            /* ParseUint("") after the sign */
    uint64_t n = 0;
    const uint64_t cutoff = UINT64_MAX / 10 + 1;           /* ParseUint: n >= cutoff -> n*10 overflows */
    for (; ok && p < len; ++p) {
      const unsigned c = s[p];
      if (c < '0' || c > '9') { ok = 0; break; }           /* syntaxError */
      if (n >= cutoff) { ok = 0; break; }                  /* rangeError */
      n *= 10;
      const uint64_t n1 = n + (c - '0');
      if (n1 < n) { ok = 0; break; }                       /* rangeError: n+d overflows */
      n = n1;
    }
    const uint64_t limit = (uint64_t)1 << 63;              /* ParseInt: cutoff := 1 << (bitSize-1) */
    if (ok && !neg && n >= limit) ok = 0;
    if (ok && neg && n > limit) ok = 0;
    values[i] = ok ? (neg ? (int64_t)(0 - n) : (int64_t)n) : 0;
    if (fail) fail[i] = ok ? 0 : KSA_FAIL_POLL;
    bad += !ok;
  }
  return bad;
}
