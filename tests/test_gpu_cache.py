"""Binary graph cache (hera_graph_save / hera_graph_load_cache): a graph restored from the cache gives the same
tables, walks and groups as one built from the PAF text -- i.e. the oracle's."""
import numpy as np
import pytest

import common
from test_gpu_parity import GRAPH_SECTIONS, GROUP_SECTIONS, PATH_SECTIONS

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", ["small_default", "mixed_sqrt", "bundled_min"])
def test_cache_round_trip(case, oracle_bin, datasets, tmp_path):
    from telomeri_b200 import capi, pipeline
    golden = common.load_golden(case)
    ds = datasets[golden["dataset"]]
    o = pipeline.parse_options(golden["options"])
    ora = common.run_dump(oracle_bin, ds, tmp_path / "o.bin", golden["options"], extra=["--rng", "philox", "--seed", "3"])
    cache = tmp_path / "graph.herakept"
    with capi.Hera(0) as h:
        pipeline.run_hot_path(ds, golden["options"], seed=3, hera=h, stages="graph")
        h.save_graph(cache)
    with capi.Hera(0) as h:
        h.load_graph_cache(cache)
        with pytest.raises(capi.HeraError):          # no PAF text on top of a restored graph
            h.load_bytes(b"r1 1 0 1 + r2 1 0 1 1 1 0\n", False)
        h.finalize(o["metrics"])
        gpu = h.export_graph()
        gpu.update(pipeline.slot_sections(gpu))
        pp = capi.default_path_params(o["rb_att"], o["bt_att"], o["len_thr"], o["w_size"], o["r_thr"])
        h.build_paths([("mc", m) for m in o["metrics"]] + [("det", m) for m in o["metrics"]], pp, 3)
        gpu.update(h.export_paths("p_"))
        h.unique()
        gpu.update({k: v for k, v in h.export_paths("u_").items() if k.startswith("u_")})
        gpu.update(h.groups(pp))
    names = [s for s in GRAPH_SECTIONS if s not in ("load_ctr", "load_cum_edges")] + PATH_SECTIONS + GROUP_SECTIONS
    common.assert_sections_equal(ora, gpu, names, case + " from cache")
    assert cache.stat().st_size < 64 + 60 * max(1, gpu["e_q"].size) + 40 * gpu["node_len"].size + 64
