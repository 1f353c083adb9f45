"""Bookkeeping of the tuned-table sweep (role of benchmark.py:227-341 + the CSV plan cache): candidate enumeration,
rank partition, checkpoint/resume with a crashing candidate, winner collection, CSV round trip. No GPU: the timer and
the checker are injected."""
import json
import os

import pytest

from deepgemm_ascend_b200 import autotune as at


def _fake_time(shape, params, layout_b):
    # deterministic pseudo-times: the 2-CTA 224-wide tile wins on the prefill shape, split 8 / 4 stages on decode
    if params["kernelSerial"] == at.KERNEL_SPLITK:
        return 20.0 - params["splitkFactor"] - (0.5 if params["stages"] == 4 else 0.0)
    return 100.0 - (10.0 if params.get("ctaGroup") == 2 else 0.0) - (5.0 if params.get("n1") == 224 else 0.0) \
        + (1.0 if params["stages"] else 0.0) + (3.0 if params["kernelSerial"] == at.KERNEL_STREAMK else 0.0)


def test_candidates_and_partition():
    c = at.candidate_plans(4096, 7168, 2048)
    assert len(c) > 20 and all(p["kernelSerial"] in (0, 4) for p in c)
    assert any(p["kernelSerial"] == 3 and p["splitkFactor"] == 8 for p in at.candidate_plans(16, 4096, 7168))
    assert all(p["n1"] % 128 == 0 for p in at.candidate_plans(512, 1024, 512, layout_b=0) if p["ctaGroup"] == 2)
    total = len(c)
    got = []
    for r in range(3):
        s, e = at.process_range(total, r, 3)
        got += list(range(s, e))
    assert got == list(range(total))          # contiguous, disjoint, complete (benchmark.py:247-253)


def test_checkpoint_resume_and_collect(tmp_path, dga):
    shape = (4096, 7168, 2048)
    cands = at.candidate_plans(*shape)
    calls = []

    def crashing_time(s, p, lb):
        calls.append(p)
        if len(calls) == 2:
            raise KeyboardInterrupt   # the process dies inside candidate #1
        return _fake_time(s, p, lb)

    run = at.AutotuneRunner([shape], 0, 1, str(tmp_path), time_fn=crashing_time, check_fn=lambda *a: 0.0)
    with pytest.raises(KeyboardInterrupt):
        run.run()
    ck = os.path.join(str(tmp_path), "shape_4096_7168_2048_rank_0_checkpoint.jsonl")
    assert json.load(open(ck))["last_process_idx"] == 1
    # restart: candidate 1 is recorded as failed (time -1) and skipped, the rest is timed, nothing is repeated
    run2 = at.AutotuneRunner([shape], 0, 1, str(tmp_path), time_fn=_fake_time, check_fn=lambda *a: 0.0)
    run2.run()
    rows = [json.loads(line) for line in open(os.path.join(str(tmp_path), "shape_4096_7168_2048_rank_0.jsonl"))]
    assert sorted(r["idx"] for r in rows) == list(range(len(cands)))
    assert [r for r in rows if r["idx"] == 1][0]["time"] == -1
    assert json.load(open(ck))["last_process_idx"] == -1
    run2.run()   # a finished sweep is a no-op
    assert len(open(os.path.join(str(tmp_path), "shape_4096_7168_2048_rank_0.jsonl")).readlines()) == len(cands)
    best = at.collect(str(tmp_path))
    t, params = best[shape]
    assert params["ctaGroup"] == 2 and params["n1"] == 224 and params["kernelSerial"] == 0 and params["stages"] == 0
    # wrong results never win
    bad = at.AutotuneRunner([(16, 4096, 7168)], 0, 1, str(tmp_path), time_fn=lambda *a: 1.0,
                            check_fn=lambda s, p, lb: 0.5 if p.get("splitkFactor") == 8 else 0.0)
    bad.run()
    assert at.collect(str(tmp_path))[(16, 4096, 7168)][1].get("splitkFactor") != 8


def test_two_ranks_then_csv_feeds_the_plan_cache(tmp_path, dga):
    shapes = [(4096, 7168, 2048), (16, 4096, 7168)]
    for r in range(2):
        at.AutotuneRunner(shapes, r, 2, str(tmp_path), time_fn=_fake_time, check_fn=lambda *a: 0.0).run()
    best = at.collect(str(tmp_path))
    assert best[(16, 4096, 7168)][1] == {"kernelSerial": 3, "splitkFactor": 8, "stages": 4}
    csv_path = str(tmp_path / "tuned.csv")
    assert at.write_csv(best, csv_path) == 2
    head = open(csv_path).readline()
    assert head.startswith("m,n,k,m1,n1,k1,kernelSerial,paddingTagA,paddingTagB,paddingTagC,blockDim")  # csv.cpp:23-26
    lib = dga._lib.lib
    lib.dgaTilingCacheClear()
    assert lib.dgaTilingCacheLoadCsv(csv_path.encode()) == 0
    p = dga.select_kernel(16, 4096, 7168, use_cache=True)
    assert (p["kernelSerial"], p["splitkFactor"], p["stages"]) == (3, 8, 4)
    p = dga.select_kernel(4096, 7168, 2048, use_cache=True)
    assert (p["ctaGroup"], p["n1"]) == (2, 224)
    lib.dgaTilingCacheClear()


def test_shipped_tuned_table_loads_and_overrides_the_selector(dga):
    """deepgemm_ascend_b200/tuned/*.csv: every row passes the library's plan validation, and a loaded row wins over
    the analytic plan for its shape (first writer wins, cache.cpp:69-77)."""
    import csv
    import os
    lib = dga._lib.lib
    path = os.path.join(os.path.dirname(dga.__file__), "tuned", "b200_deepseek_v3_bf16.csv")
    rows = list(csv.DictReader(open(path)))
    assert rows
    lib.dgaTilingCacheClear()
    try:
        assert lib.dgaTilingCacheLoadCsv(path.encode()) == 0
        assert lib.dgaTilingCacheSize() == len(rows)
        for r in rows:
            got = dga.select_kernel(int(r["m"]), int(r["n"]), int(r["k"]), use_cache=True)
            assert (got["n1"], got["ctaGroup"], got["kernelSerial"]) == (int(r["n1"]), int(r["ctaGroup"]), int(r["kernelSerial"]))
            assert got["smemBytes"] <= 232448
    finally:
        lib.dgaTilingCacheClear()
