"""A stream of independent adjustment batches through two handles (host logic over include/viso_ba.h).

The reference's adjuster holds ONE g2o graph for one window (its constructor, reference
include/bundle_adjuster.h:111-166) and has no notion of a stream of windows.  A back end that is handed
batch after batch of windows (BASELINE configs[4]) has three phases per batch: host->device + structure build (PCIe and short
kernels), the LM loop (fills the GPU), device->host.  One handle = one CUDA stream, so two handles on two
host threads let a batch's transfers hide behind the other batch's LM loop -- provided the two threads do
not fall into step (both copying, then both iterating: measured, tools/e2e_phases.py).  The LM loops
therefore take turns under a lock; uploads and read-backs stay outside it.  ctypes releases the GIL
inside every C ABI call.  Results are bit-identical to calling Optimize on one handle.
"""
import threading


def optimize_stream(handles, batches, iterations, consume=None):
    """Run every batch of `batches` (dicts with the arrays of api.batch_problems, or of a single problem:
    poses_init, points_init, edge_pose, edge_point, edge_uv [, pose_off, point_off, edge_off]) through
    `handles` (BundleAdjuster objects of sufficient capacity, normally two).  For batch i, in the worker that
    owns it: set_problem* -> upload -> [lock] optimize_resident [/lock] -> download -> consume(i, handle,
    summary).  `consume` reads what it needs from the handle (poses(), points(), problem_summaries()) before
    the handle is re-used.  Returns the summaries in batch order."""
    batches = list(batches)
    out = [None] * len(batches)
    errors = []
    lm_turn = threading.Lock()

    def load(ba, b):
        if "pose_off" in b and len(b["pose_off"]) > 2:
            ba.set_problem_batched(b["pose_off"], b["point_off"], b["edge_off"], b["poses_init"], b["points_init"],
                                   b["edge_pose"], b["edge_point"], b["edge_uv"])
        else:
            ba.set_problem(b["poses_init"], b["points_init"], b["edge_pose"], b["edge_point"], b["edge_uv"])

    def worker(j):
        try:
            ba = handles[j]
            for i in range(j, len(batches), len(handles)):
                load(ba, batches[i])
                ba.upload()
                with lm_turn:
                    s = ba.optimize_resident(iterations)
                ba.download()
                out[i] = s
                if consume is not None:
                    consume(i, ba, s)
        except Exception as e:  # surfaced to the caller below
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(j,)) for j in range(len(handles))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
    return out
