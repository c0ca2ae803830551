"""SURVEY.md 8(b) threading contract: the callers are rayon workers and tokio blocking threads, several STARK proofs
may commit at once (starkyx machine/emulated/stark.rs:190-200), so every entry point must be callable from many host
threads -- on one shared context (calls serialise on it) and on one context per thread (calls run concurrently on
their own streams).  Every result is checked bit-exact against the oracle.
"""
import gc
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SHAPES = [(9, 10, 1, 4), (33, 9, 3, 4), (5, 12, 1, 2), (40, 8, 2, 3), (17, 11, 1, 4), (64, 9, 1, 4)]


def _jobs(O, rng, copies):
    jobs = []
    for k in range(copies):
        for W, log_n, r, cap in SHAPES:
            v = O.random_field(rng, (W, 1 << log_n))
            jobs.append((v, r, cap, O.commit_from_values(v, r, cap)))
    return jobs


def _run(threads):
    errs = []

    def wrap(fn):
        def inner():
            try:
                fn()
            except BaseException as e:  # noqa: BLE001 -- reported to the main thread
                errs.append(repr(e))
        return inner
    ts = [threading.Thread(target=wrap(f)) for f in threads]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=300)
    assert not any(t.is_alive() for t in ts), "a worker thread hung"
    assert not errs, errs


def _commit_and_check(P2, O, ctx, job):
    v, r, cap, ref = job
    b = P2.PolynomialBatch.from_values([P2.PolynomialValues(c) for c in v], r, False, cap, None, None, ctx=ctx)
    assert np.array_equal(b.merkle_tree.cap, ref.cap)
    assert np.array_equal(b.coeff_matrix, ref.coeffs)
    i = (v.shape[1] * 3 + 1) % (v.shape[1] << r)
    row, proof = b.merkle_tree.get(i), b.merkle_tree.prove(i)
    assert np.array_equal(row, ref.leaves[i])
    assert O.merkle_verify(row, i, ref.cap, np.asarray(proof.siblings, np.uint64))


def test_shared_context_from_many_threads(P2, O, rng):
    jobs = _jobs(O, rng, 2)
    ctx = P2.default_context()
    n_threads = 6
    _run([lambda t=t: [_commit_and_check(P2, O, ctx, j) for j in jobs[t::n_threads]] for t in range(n_threads)])


def test_one_context_per_thread(P2, O, rng):
    jobs = _jobs(O, rng, 2)
    n_threads = 4
    ctxs = [P2.Context(0) for _ in range(n_threads)]
    try:
        _run([lambda t=t: [_commit_and_check(P2, O, ctxs[t], j) for j in jobs[t::n_threads]] for t in range(n_threads)])
    finally:
        gc.collect()  # batches sit in a reference cycle with their merkle_tree view: free them before their context
        for c in ctxs:
            c.close()


def test_errors_on_one_thread_do_not_disturb_another(P2, O, rng):
    jobs = _jobs(O, rng, 1)
    ctx = P2.default_context()

    def bad():
        for _ in range(20):
            with pytest.raises(P2.NlcError):
                P2.PolynomialBatch.from_values([P2.PolynomialValues(np.ones(8, np.uint64))], 1, False, 9, ctx=ctx)

    _run([bad, lambda: [_commit_and_check(P2, O, ctx, j) for j in jobs]])
