// Pipeline-schedule generator, validator and discrete-event simulator (C ABI in
// include/sdml_schedule.h).  Replaces the reference's implicit schedule -- one batch in flight,
// rank 1 idle while rank 0 computes and vice versa (/root/reference/simple_distributed.py:108-113,
// the blocking rpc_sync at ref:49) -- which is what "lockstep" reproduces.
//
// gpipe / 1f1b / lockstep are closed-form.  The zero-bubble variants (zb-h1, zb-h2)
// are built by a causal list scheduler: it walks simulated time, and on every rank
// picks, among the ops whose inputs have arrived, B first (it unblocks the upstream
// rank), then F (if the in-flight activation budget allows), and fills any idle gap
// with pending W's.  The in-flight budget is what distinguishes h1 (p - rank, the
// 1F1B footprint) from h2 (2 (p - rank) - 1).
#include "sdml_schedule.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

namespace {

struct Op {
  int kind;
  int mb;
};
using RankOps = std::vector<Op>;

constexpr double kInf = std::numeric_limits<double>::infinity();

bool known_kind(const std::string& k) {
  return k == "gpipe" || k == "1f1b" || k == "zb-h1" || k == "zb-h2" || k == "lockstep";
}

std::vector<RankOps> gen_gpipe(int p, int m) {
  std::vector<RankOps> out(p);
  for (int s = 0; s < p; ++s) {
    for (int i = 0; i < m; ++i) out[s].push_back({SDML_OP_F, i});
    for (int i = 0; i < m; ++i) out[s].push_back({SDML_OP_BW, i});
  }
  return out;
}

std::vector<RankOps> gen_1f1b(int p, int m) {
  std::vector<RankOps> out(p);
  for (int s = 0; s < p; ++s) {
    const int warm = std::min(p - 1 - s, m);
    for (int i = 0; i < warm; ++i) out[s].push_back({SDML_OP_F, i});
    for (int i = 0; i < m - warm; ++i) {
      out[s].push_back({SDML_OP_F, warm + i});
      out[s].push_back({SDML_OP_BW, i});
    }
    for (int i = m - warm; i < m; ++i) out[s].push_back({SDML_OP_BW, i});
  }
  return out;
}

// Causal greedy construction with split B / W.
std::vector<RankOps> gen_zb(int p, int m, const std::vector<int>& limit, double tf, double tb,
                            double tw, double tc) {
  std::vector<RankOps> out(p);
  std::vector<double> t_free(p, 0.0);
  std::vector<int> next_f(p, 0), next_b(p, 0), next_w(p, 0);
  // f_end[s][i], b_end[s][i]: completion times (inf = not scheduled yet)
  std::vector<std::vector<double>> f_end(p, std::vector<double>(m, kInf));
  std::vector<std::vector<double>> b_end(p, std::vector<double>(m, kInf));
  const int total = 3 * p * m;
  for (int done = 0; done < total; ++done) {
    int best_rank = -1, best_kind = -1;
    double best_start = kInf;
    for (int s = 0; s < p; ++s) {
      double sb = kInf, sf = kInf;
      if (next_b[s] < m && next_b[s] < next_f[s]) {
        const int i = next_b[s];
        const double arr = (s == p - 1) ? f_end[s][i] : b_end[s + 1][i] + tc;
        if (arr < kInf) sb = std::max(t_free[s], arr);
      }
      if (next_f[s] < m && next_f[s] - next_b[s] < limit[s]) {
        const int i = next_f[s];
        const double arr = (s == 0) ? 0.0 : f_end[s - 1][i] + tc;
        if (arr < kInf) sf = std::max(t_free[s], arr);
      }
      const bool w_pending = next_w[s] < next_b[s];
      int kind = -1;
      double start = kInf;
      const double eps = 1e-12;
      if (sb <= t_free[s] + eps) { kind = SDML_OP_B; start = sb; }
      else if (sf <= t_free[s] + eps) { kind = SDML_OP_F; start = sf; }
      else if (w_pending) { kind = SDML_OP_W; start = t_free[s]; }
      else if (sb <= sf && sb < kInf) { kind = SDML_OP_B; start = sb; }
      else if (sf < kInf) { kind = SDML_OP_F; start = sf; }
      if (kind >= 0 && start < best_start - 1e-12) {
        best_start = start; best_rank = s; best_kind = kind;
      }
    }
    if (best_rank < 0) return {};  // should not happen: dead-locked construction
    const int s = best_rank;
    if (best_kind == SDML_OP_F) {
      const int i = next_f[s]++;
      f_end[s][i] = best_start + tf;
      t_free[s] = f_end[s][i];
      out[s].push_back({SDML_OP_F, i});
    } else if (best_kind == SDML_OP_B) {
      const int i = next_b[s]++;
      b_end[s][i] = best_start + tb;
      t_free[s] = b_end[s][i];
      out[s].push_back({SDML_OP_B, i});
    } else {
      const int i = next_w[s]++;
      t_free[s] = best_start + tw;
      out[s].push_back({SDML_OP_W, i});
    }
  }
  return out;
}

int build(const std::string& kind, int p, int m, double tf, double tb, double tw, double tc,
          std::vector<RankOps>& out) {
  if (!known_kind(kind)) return -1;
  if (p < 1 || m < 1) return -2;
  if (kind == "lockstep") {
    if (m != 1) return -3;
    out = gen_gpipe(p, 1);
  } else if (kind == "gpipe") {
    out = gen_gpipe(p, m);
  } else if (kind == "1f1b") {
    out = gen_1f1b(p, m);
  } else {
    std::vector<int> limit(p);
    for (int s = 0; s < p; ++s)
      limit[s] = std::max(1, kind == "zb-h1" ? (p - s) : (2 * (p - s) - 1));
    if (!(tf > 0) || !(tb > 0) || !(tw > 0) || tc < 0) { tf = tb = tw = 1.0; tc = 0.0; }
    out = gen_zb(p, m, limit, tf, tb, tw, tc);
    if (out.empty()) return -4;
  }
  return 0;
}

// Event simulation of an arbitrary per-rank op list.
int simulate(const std::vector<RankOps>& ops, int p, int m, double tf, double tb, double tw,
             double tc, double* makespan, double* busy, int* peak) {
  std::vector<size_t> pc(p, 0);
  std::vector<double> t_free(p, 0.0), busy_v(p, 0.0);
  std::vector<int> inflight(p, 0), peak_v(p, 0);
  std::vector<std::vector<double>> f_end(p, std::vector<double>(m, kInf));
  std::vector<std::vector<double>> b_end(p, std::vector<double>(m, kInf));
  size_t remaining = 0;
  for (auto& r : ops) remaining += r.size();
  double first_start = kInf, last_end = 0.0;
  while (remaining > 0) {
    bool progressed = false;
    for (int s = 0; s < p; ++s) {
      while (pc[s] < ops[s].size()) {
        const Op op = ops[s][pc[s]];
        double arr = 0.0, dur = 0.0;
        if (op.kind == SDML_OP_F) {
          arr = (s == 0) ? 0.0 : f_end[s - 1][op.mb] + tc;
          dur = tf;
        } else if (op.kind == SDML_OP_B || op.kind == SDML_OP_BW) {
          arr = (s == p - 1) ? f_end[s][op.mb] : b_end[s + 1][op.mb] + tc;
          dur = op.kind == SDML_OP_B ? tb : tb + tw;
        } else {
          arr = b_end[s][op.mb];
          dur = tw;
        }
        if (!(arr < kInf)) break;  // producer not simulated yet
        const double start = std::max(t_free[s], arr);
        const double end = start + dur;
        if (op.kind == SDML_OP_F) {
          f_end[s][op.mb] = end;
          peak_v[s] = std::max(peak_v[s], ++inflight[s]);
        } else if (op.kind == SDML_OP_B) {
          b_end[s][op.mb] = end;
          --inflight[s];
        } else if (op.kind == SDML_OP_BW) {
          b_end[s][op.mb] = start + tb;  // dX leaves after the dgrad half
          --inflight[s];
        }
        t_free[s] = end;
        busy_v[s] += dur;
        first_start = std::min(first_start, start);
        last_end = std::max(last_end, end);
        ++pc[s];
        --remaining;
        progressed = true;
      }
    }
    if (!progressed) return -5;  // dead-lock
  }
  if (makespan) *makespan = last_end - first_start;
  if (busy) std::copy(busy_v.begin(), busy_v.end(), busy);
  if (peak) std::copy(peak_v.begin(), peak_v.end(), peak);
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Interleaved (virtual-stage) zero-bubble schedules: every rank holds v model chunks; virtual
// stage q = c * p + r is chunk c of rank r, so a micro-batch crosses the ranks v times on its way
// down and v times on its way back.  The ramp a rank waits for shrinks from (p - 1) stage-sized
// forwards to (p - 1) CHUNK-sized ones: with unit costs per stage the single-step bubble bound
// becomes ((p - 1) / v) / (3m + (p - 1) / v)  (p = 4, v = 2, m = 8: 5.9 % instead of 11.1 %).
// Same causal list scheduler as gen_zb: B first (it unblocks the upstream virtual stage), then F
// (deepest ready chunk first, so a micro-batch reaches the loss as early as possible) within an
// in-flight budget, pending W's fill idle time.
// ---------------------------------------------------------------------------------------------
struct OpV {
  int kind, mb, chunk;
};
using RankOpsV = std::vector<OpV>;

struct VState {
  int p, v, m;
  double tf, tb, tw, tc;
  // [q][i] completion times of F / B of virtual stage q, micro-batch i
  std::vector<std::vector<double>> f_end, b_end;
  VState(int p_, int v_, int m_, double tf_, double tb_, double tw_, double tc_)
      : p(p_), v(v_), m(m_), tf(tf_), tb(tb_), tw(tw_), tc(tc_),
        f_end(p_ * v_, std::vector<double>(m_, kInf)), b_end(p_ * v_, std::vector<double>(m_, kInf)) {}
  int q_of(int rank, int chunk) const { return chunk * p + rank; }
  double f_arrival(int q, int i) const { return q == 0 ? 0.0 : f_end[q - 1][i] + tc; }
  double b_arrival(int q, int i) const { return q == p * v - 1 ? f_end[q][i] : b_end[q + 1][i] + tc; }
};

std::vector<RankOpsV> gen_zb_interleaved(int p, int v, int m, int limit_chunks, double tf, double tb,
                                         double tw, double tc) {
  VState st(p, v, m, tf, tb, tw, tc);
  std::vector<RankOpsV> out(p);
  std::vector<double> t_free(p, 0.0);
  // per (rank, chunk): next micro-batch to forward / backward / weight-grad
  std::vector<std::vector<int>> nf(p, std::vector<int>(v, 0)), nb(p, std::vector<int>(v, 0)),
      nw(p, std::vector<int>(v, 0));
  std::vector<int> inflight(p, 0);
  const long long total = 3LL * p * v * m;
  const double eps = 1e-12;
  for (long long done = 0; done < total; ++done) {
    int best_rank = -1, best_kind = -1, best_chunk = -1;
    double best_start = kInf;
    for (int s = 0; s < p; ++s) {
      // earliest-ready B over the chunks (ties: the deeper chunk, its gradient has further to go)
      double sb = kInf, sf = kInf;
      int cb = -1, cf = -1, cw = -1;
      for (int c = v - 1; c >= 0; --c) {
        const int q = st.q_of(s, c);
        if (nb[s][c] < m && nb[s][c] < nf[s][c]) {
          const double arr = st.b_arrival(q, nb[s][c]);
          if (arr < kInf) {
            const double t = std::max(t_free[s], arr);
            if (t < sb - eps) { sb = t; cb = c; }
          }
        }
        if (nf[s][c] < m && inflight[s] < limit_chunks) {
          const double arr = st.f_arrival(q, nf[s][c]);
          if (arr < kInf) {
            const double t = std::max(t_free[s], arr);
            if (t < sf - eps) { sf = t; cf = c; }   // scanning from the deepest chunk: it wins ties
          }
        }
        if (cw < 0 && nw[s][c] < nb[s][c]) cw = c;
      }
      int kind = -1, chunk = -1;
      double start = kInf;
      if (cb >= 0 && sb <= t_free[s] + eps) { kind = SDML_OP_B; chunk = cb; start = sb; }
      else if (cf >= 0 && sf <= t_free[s] + eps) { kind = SDML_OP_F; chunk = cf; start = sf; }
      else if (cw >= 0) { kind = SDML_OP_W; chunk = cw; start = t_free[s]; }
      else if (cb >= 0 && sb <= sf) { kind = SDML_OP_B; chunk = cb; start = sb; }
      else if (cf >= 0) { kind = SDML_OP_F; chunk = cf; start = sf; }
      else if (cb >= 0) { kind = SDML_OP_B; chunk = cb; start = sb; }
      if (kind >= 0 && start < best_start - eps) {
        best_start = start; best_rank = s; best_kind = kind; best_chunk = chunk;
      }
    }
    if (best_rank < 0) return {};
    const int s = best_rank, c = best_chunk, q = st.q_of(s, c);
    if (best_kind == SDML_OP_F) {
      const int i = nf[s][c]++;
      st.f_end[q][i] = best_start + tf;
      t_free[s] = st.f_end[q][i];
      ++inflight[s];
      out[s].push_back({SDML_OP_F, i, c});
    } else if (best_kind == SDML_OP_B) {
      const int i = nb[s][c]++;
      st.b_end[q][i] = best_start + tb;
      t_free[s] = st.b_end[q][i];
      --inflight[s];
      out[s].push_back({SDML_OP_B, i, c});
    } else {
      const int i = nw[s][c]++;
      t_free[s] = best_start + tw;
      out[s].push_back({SDML_OP_W, i, c});
    }
  }
  return out;
}

int simulate_interleaved(const std::vector<RankOpsV>& ops, int p, int v, int m, double tf, double tb,
                         double tw, double tc, double* makespan, double* busy, int* peak) {
  VState st(p, v, m, tf, tb, tw, tc);
  std::vector<size_t> pc(p, 0);
  std::vector<double> t_free(p, 0.0), busy_v(p, 0.0);
  std::vector<int> inflight(p, 0), peak_v(p, 0);
  size_t remaining = 0;
  for (auto& r : ops) remaining += r.size();
  double first_start = kInf, last_end = 0.0;
  while (remaining > 0) {
    bool progressed = false;
    for (int s = 0; s < p; ++s) {
      while (pc[s] < ops[s].size()) {
        const OpV op = ops[s][pc[s]];
        const int q = st.q_of(s, op.chunk);
        double arr, dur;
        if (op.kind == SDML_OP_F) { arr = st.f_arrival(q, op.mb); dur = tf; }
        else if (op.kind == SDML_OP_B) { arr = st.b_arrival(q, op.mb); dur = tb; }
        else { arr = st.b_end[q][op.mb]; dur = tw; }
        if (!(arr < kInf)) break;
        const double start = std::max(t_free[s], arr), end = start + dur;
        if (op.kind == SDML_OP_F) { st.f_end[q][op.mb] = end; peak_v[s] = std::max(peak_v[s], ++inflight[s]); }
        else if (op.kind == SDML_OP_B) { st.b_end[q][op.mb] = end; --inflight[s]; }
        t_free[s] = end;
        busy_v[s] += dur;
        first_start = std::min(first_start, start);
        last_end = std::max(last_end, end);
        ++pc[s]; --remaining; progressed = true;
      }
    }
    if (!progressed) return -5;
  }
  if (makespan) *makespan = last_end - first_start;
  if (busy) std::copy(busy_v.begin(), busy_v.end(), busy);
  if (peak) std::copy(peak_v.begin(), peak_v.end(), peak);
  return 0;
}

int build_interleaved(int p, int v, int m, int limit_chunks, double& tf, double& tb, double& tw, double& tc,
                      std::vector<RankOpsV>& out) {
  if (p < 1 || v < 1 || m < 1) return -2;
  if (!(tf > 0) || !(tb > 0) || !(tw > 0) || tc < 0) { tf = tb = tw = 1.0; tc = 0.0; }
  if (limit_chunks <= 0) limit_chunks = 2 * p * v;   // generous: memory is not the constraint on 180 GB parts
  out = gen_zb_interleaved(p, v, m, limit_chunks, tf, tb, tw, tc);
  return out.empty() ? -4 : 0;
}

}  // namespace

// Interleaved zero-bubble schedule: v chunks per rank; ops are (kind, micro-batch, chunk) triples.
// Costs are PER CHUNK op.  limit_chunks: max forwards in flight on a rank (0 = 2 * p * v).
extern "C" int sdml_sched_generate_interleaved(int p, int v, int m, int rank, int limit_chunks, double t_f,
                                               double t_b, double t_w, double t_c, int* out_ops, int max_ops) {
  if (!out_ops || rank < 0 || rank >= p) return -2;
  std::vector<RankOpsV> all;
  const int rc = build_interleaved(p, v, m, limit_chunks, t_f, t_b, t_w, t_c, all);
  if (rc) return rc;
  const RankOpsV& r = all[rank];
  if (static_cast<int>(r.size()) > max_ops) return -6;
  for (size_t i = 0; i < r.size(); ++i) {
    out_ops[3 * i] = r[i].kind;
    out_ops[3 * i + 1] = r[i].mb;
    out_ops[3 * i + 2] = r[i].chunk;
  }
  return static_cast<int>(r.size());
}

extern "C" int sdml_sched_simulate_interleaved(int p, int v, int m, int limit_chunks, double t_f, double t_b,
                                               double t_w, double t_c, double* makespan, double* busy,
                                               int* peak_inflight) {
  std::vector<RankOpsV> all;
  const int rc = build_interleaved(p, v, m, limit_chunks, t_f, t_b, t_w, t_c, all);
  if (rc) return rc;
  return simulate_interleaved(all, p, v, m, t_f, t_b, t_w, t_c, makespan, busy, peak_inflight);
}

extern "C" int sdml_sched_generate(const char* kind, int p, int m, int rank, double t_f,
                                   double t_b, double t_w, double t_c, int* out_ops,
                                   int max_ops) {
  if (!kind || !out_ops) return -1;
  if (rank < 0 || rank >= p) return -2;
  std::vector<RankOps> all;
  const int rc = build(kind, p, m, t_f, t_b, t_w, t_c, all);
  if (rc) return rc;
  const RankOps& r = all[rank];
  if (static_cast<int>(r.size()) > max_ops) return -6;
  for (size_t i = 0; i < r.size(); ++i) {
    out_ops[2 * i] = r[i].kind;
    out_ops[2 * i + 1] = r[i].mb;
  }
  return static_cast<int>(r.size());
}

extern "C" int sdml_sched_simulate(const char* kind, int p, int m, double t_f, double t_b,
                                   double t_w, double t_c, double* makespan, double* busy,
                                   int* peak_inflight) {
  if (!kind) return -1;
  std::vector<RankOps> all;
  const int rc = build(kind, p, m, t_f, t_b, t_w, t_c, all);
  if (rc) return rc;
  return simulate(all, p, m, t_f, t_b, t_w, t_c, makespan, busy, peak_inflight);
}

extern "C" int sdml_sched_validate(const int* ops, int n_ops, int m) {
  if (!ops || m < 1) return -1;
  std::vector<int> f_pos(m, -1), b_pos(m, -1), w_pos(m, -1);
  int last_f = -1, last_b = -1;
  bool split = false;
  for (int i = 0; i < n_ops; ++i) {
    const int kind = ops[2 * i], mb = ops[2 * i + 1];
    if (mb < 0 || mb >= m) return -2;
    switch (kind) {
      case SDML_OP_F:
        if (f_pos[mb] >= 0) return -3;
        if (mb != last_f + 1) return -4;
        last_f = mb; f_pos[mb] = i;
        break;
      case SDML_OP_B:
        split = true;
        // fallthrough
      case SDML_OP_BW:
        if (b_pos[mb] >= 0) return -5;
        if (mb != last_b + 1) return -6;
        last_b = mb; b_pos[mb] = i;
        if (kind == SDML_OP_BW) w_pos[mb] = i;
        break;
      case SDML_OP_W:
        if (w_pos[mb] >= 0) return -7;
        w_pos[mb] = i;
        break;
      default:
        return -8;
    }
  }
  for (int i = 0; i < m; ++i) {
    if (f_pos[i] < 0 || b_pos[i] < 0 || w_pos[i] < 0) return -9;
    if (!(f_pos[i] < b_pos[i] && b_pos[i] <= w_pos[i])) return -10;
  }
  (void)split;
  return 0;
}
