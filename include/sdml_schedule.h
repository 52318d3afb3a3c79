/* C ABI of libsdml_sched.so -- pipeline-schedule generator, validator and simulator
 * (csrc/schedule.cpp).  Replaces the reference's lock-step, one-batch-in-flight loop
 * (/root/reference/simple_distributed.py:108-113: forward -> rpc_sync -> backward ->
 * step, rank 1 idle while rank 0 computes and vice versa) with micro-batched
 * schedules.
 *
 * A schedule is, per pipeline rank, an ordered list of ops (kind, micro-batch):
 *   SDML_OP_F   forward of one micro-batch through this rank's stage
 *   SDML_OP_B   input-gradient part of the backward (dgrad; sends dX upstream)
 *   SDML_OP_W   weight-gradient part of the backward (wgrad; no communication)
 *   SDML_OP_BW  fused backward (B immediately followed by W)
 * Receives are implied: F(mb) on rank s>0 consumes the activation produced by F(mb) on
 * rank s-1; B/BW(mb) on rank s<p-1 consumes the gradient produced by B/BW(mb) on s+1.
 * Every generator issues F's in ascending micro-batch order and B's in ascending
 * micro-batch order on every rank, so send/recv pairs match by order.
 *
 * Schedules: "gpipe", "1f1b", "zb-h1" (B/W split, 1F1B memory), "zb-h2" (B/W split,
 * ~2x in-flight activations, bubble-free when F=B=W), "lockstep" (the reference's
 * behaviour: m must be 1).
 */
#ifndef SDML_SCHEDULE_H_
#define SDML_SCHEDULE_H_

#ifdef __cplusplus
extern "C" {
#endif

enum { SDML_OP_F = 0, SDML_OP_B = 1, SDML_OP_W = 2, SDML_OP_BW = 3 };

/* Writes up to max_ops (kind, mb) pairs for `rank` into out_ops[2*i], out_ops[2*i+1].
 * Returns the number of ops, or a negative value on error (unknown kind, bad sizes,
 * buffer too small).  t_f/t_b/t_w/t_c are the relative costs used by the zero-bubble
 * heuristic (forward, input-grad, weight-grad, stage-to-stage transfer); pass 1,1,1,0
 * when unknown. */
int sdml_sched_generate(const char* kind, int p, int m, int rank, double t_f, double t_b,
                        double t_w, double t_c, int* out_ops, int max_ops);

/* Simulates the whole pipeline under the given costs.  Outputs (any may be NULL):
 *   makespan       : time from first op start to last op end
 *   busy[p]        : summed op time per rank
 *   peak_inflight[p]: max number of micro-batches with F done and B not done
 * Returns 0, or negative when the schedule deadlocks / is invalid. */
int sdml_sched_simulate(const char* kind, int p, int m, double t_f, double t_b, double t_w,
                        double t_c, double* makespan, double* busy, int* peak_inflight);

/* Structural check of a per-rank op list (as produced by generate): every micro-batch
 * has exactly one F, one B-or-BW and (when B is split) one W, in F < B <= W order, and
 * F's / B's ascend.  Returns 0 when valid, else a negative code. */
int sdml_sched_validate(const int* ops, int n_ops, int m);

/* Interleaved (virtual-stage) zero-bubble schedules: every rank holds v model chunks, virtual stage
 * c * p + r is chunk c of rank r.  Ops are (kind, micro-batch, chunk) TRIPLES (out_ops[3*i .. 3*i+2]);
 * kinds F / B / W only.  Costs are per chunk op; limit_chunks bounds the forwards in flight on a rank
 * (0 = 2 * p * v).  With unit stage costs the single-step bubble bound drops from (p-1)/(3m+p-1) to
 * ((p-1)/v) / (3m + (p-1)/v).  No counterpart in the reference (one batch in flight, ref:108-113). */
int sdml_sched_generate_interleaved(int p, int v, int m, int rank, int limit_chunks, double t_f, double t_b,
                                    double t_w, double t_c, int* out_ops, int max_ops);
int sdml_sched_simulate_interleaved(int p, int v, int m, int limit_chunks, double t_f, double t_b, double t_w,
                                    double t_c, double* makespan, double* busy, int* peak_inflight);

#ifdef __cplusplus
}
#endif
#endif /* SDML_SCHEDULE_H_ */
