"""Comparison of selection results with the reference-generated fixtures (tests/golden/*.npz), honouring the
per-row ambiguity masks the fixtures store instead of walking seeds (tests/golden/make_golden.py):

  * row not flagged          -> bit equality with the reference golden, order included;
  * `amb_rows` only          -> the reference's fp32 GEMV ordered two near-tied clusters differently from exact
                                arithmetic: the selected SET must still be the golden's;
  * `amb_set`                -> the near-tie straddles the cut: the sets may differ, but only by clusters whose
                                exact float64 score lies within the fp32 rounding band of the k-th score.
Returns / prints the number of flagged rows, as SURVEY section 7 prescribes."""
import torch


def check_indices(got, f, what=""):
    got = got.cpu().to(torch.int64)
    ref = f["ref_indices"].to(torch.int64)
    assert tuple(got.shape) == tuple(ref.shape)
    amb = f["amb_rows"].bool() if "amb_rows" in f else torch.zeros(ref.shape[:2], dtype=torch.bool)
    ambs = f["amb_set"].bool() if "amb_set" in f else torch.zeros(ref.shape[:2], dtype=torch.bool)
    clean = ~amb
    assert torch.equal(got[clean], ref[clean]), f"{what}: indices differ from the reference golden on unflagged rows"
    order_only = amb & ~ambs
    if order_only.any():
        assert torch.equal(torch.sort(got[order_only], dim=-1)[0], torch.sort(ref[order_only], dim=-1)[0]), \
            f"{what}: selected set differs on a row where only the golden's ORDER is ambiguous"
    if ambs.any():
        q, l1 = f["q"].double(), f["l1"].double()
        g = q.shape[1] // l1.shape[0]
        for b, h in ambs.nonzero().tolist():
            s = l1[h // g] @ q[b, h]                                    # exact scores [C]
            k = ref.shape[2]
            kth = torch.sort(s, descending=True)[0][k - 1]
            band = 2.0 ** -15 * q[b, h].abs().sum() * l1[h // g].abs().max()
            diff = set(got[b, h].tolist()) ^ set(ref[b, h].tolist())
            assert all(abs(float(s[c] - kth)) <= float(band) for c in diff), \
                f"{what}: row ({b},{h}) differs from the golden by a cluster that is not a near-tie of the cut"
    n = int(amb.sum())
    print(f"[golden] {what}: {n} of {amb.numel()} rows flagged ambiguous in the reference's fp32 order "
          f"({int(ambs.sum())} of them in the selected set)")
    return n


def rows_with_defined_positions(f, Bs):
    """[Bs, Nq] bool: rows whose page_position is defined by the golden (selected SET unambiguous)."""
    Nq = f["ref_indices"].shape[1]
    if "amb_set" not in f:
        return torch.ones(Bs, Nq, dtype=torch.bool)
    return ~f["amb_set"].bool()[:Bs]
