"""hsq_align_scores_noalloc (no counterpart in the reference): the DP of hsq_align_batch
with its scratch in a fixed file-scope __device__ array, for callers that cannot call cudaMalloc.  Called through the C ABI twice in a
row (the array is shared between calls and keeps the first call's contents) and compared, bit for bit, with hsq_align_batch.
This file is collected last on purpose."""
import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from tests import helpers as H

pytestmark = pytest.mark.gpu


def _device_inputs(torch, batch):
    data = torch.from_numpy(np.ascontiguousarray(batch.data) if len(batch.data) else np.zeros(1, dtype=np.uint8)).cuda()
    off = torch.from_numpy(np.ascontiguousarray(batch.offsets)).cuda()
    return data, off


def _run(ctx, torch, refs, qrys, flags, max_ref, max_qry):
    n = refs.n
    rd, ro = _device_inputs(torch, refs)
    qd, qo = _device_inputs(torch, qrys)
    outs = [torch.full((max(n, 1),), -777, dtype=torch.int32, device="cuda") for _ in range(4)]
    status = torch.full((1,), 55, dtype=torch.int32, device="cuda")  # the call must clear it
    torch.cuda.synchronize()
    ctx.align_scores_noalloc(rd.data_ptr(), ro.data_ptr(), len(refs.data), qd.data_ptr(), qo.data_ptr(), len(qrys.data), n, max_ref, max_qry,
                             *[t.data_ptr() for t in outs], status.data_ptr(), flags=flags)
    return [t.cpu().numpy()[:n] for t in outs], int(status.item())


def test_noalloc_entry_point_equals_align_batch_two_calls_in_a_row():
    import torch
    from bioinformatics_b200._lib import Context, HsqError
    rng = np.random.default_rng(31)
    with Context(0) as ctx:
        # two different batches back to back: long sequences first, so the second call runs over scratch the first one filled
        cases = []
        r1, q1 = H.align_cases(rng, 3000, max_len=150)
        cases.append((r1, q1, 0))
        r2, q2 = H.align_cases(rng, 9000, max_len=60)  # more items than resident threads (8192): the grid-stride loop is exercised
        cases.append((r2, q2, 3))
        cases.append((r1, q1, 0))  # and the first batch again: same answers as before
        for refs_s, qrys_s, flags in cases:
            refs, qrys = A.SeqBatch.from_strings(refs_s), A.SeqBatch.from_strings(qrys_s)
            max_ref, max_qry = int(refs.lengths().max()), int(qrys.lengths().max())
            (score, n_cols, matches, mismatches), status = _run(ctx, torch, refs, qrys, flags, max_ref, max_qry)
            want = ctx.align_batch(refs, qrys, flags=flags, cols_cap=400, md_cap=600)  # the existing entry point
            assert status == 0
            for name, got in (("score", score), ("n_cols", n_cols), ("matches", matches), ("mismatches", mismatches)):
                assert np.array_equal(got, getattr(want, name)), name  # integer results: bit-exact
        # limits are refused with a clear error, nothing is launched
        refs, qrys = A.SeqBatch.from_strings(["A" * 300]), A.SeqBatch.from_strings(["A" * 300])
        with pytest.raises(HsqError) as e:
            _run(ctx, torch, refs, qrys, 0, 300, 300)
        assert e.value.status == A.HSQ_ERR_UNSUPPORTED and "1792" in str(e.value)
        # a sequence longer than declared and a bad character are reported through the status word
        refs, qrys = A.SeqBatch.from_strings(["ACGTACGT", "ACGT"]), A.SeqBatch.from_strings(["ACGT", "AC!T"])
        _, status = _run(ctx, torch, refs, qrys, 0, 8, 4)
        assert status == A.HSQ_ERR_BAD_CHARACTER
        _, status = _run(ctx, torch, refs, qrys, 0, 4, 4)
        assert status in (A.HSQ_ERR_BAD_CHARACTER, A.HSQ_ERR_INVALID_ARGUMENT)
        (score, _, _, _), status = _run(ctx, torch, A.SeqBatch.from_strings([]), A.SeqBatch.from_strings([]), 0, 0, 0)
        assert status == 0 and len(score) == 0
