"""Host-side model of the SafeKV reuse rule used by the cfg-5 workload generator
(tools/bench_safekv_multiturn.py): private radix nodes are reusable only by their owner
(reference: python/sglang/srt/mem_cache/radix_cache.py:333-349)."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tool():
    spec = importlib.util.spec_from_file_location(
        "bench_safekv_multiturn", os.path.join(ROOT, "tools", "bench_safekv_multiturn.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_private_nodes_only_match_their_owner():
    tree = _tool().SafeKVPrefixModel()
    a = ["sys", "u0t0", "u0t1"]
    assert tree.match(a, 0) == 0
    tree.insert(a, 0, private_from=1)
    assert tree.match(a, 0) == 3                      # owner: shared system prompt + own history
    assert tree.match(a, 1) == 1                      # another user, same tokens: system prompt only
    assert tree.match(["sys", "u1t0"], 1) == 1        # another conversation shares the public node
    tree.insert(["sys", "u1t0"], 1, private_from=1)
    assert tree.match(["sys", "u1t0"], 1) == 2 and tree.match(["sys", "u1t0"], 0) == 1
    # matching stops at the first refused node even if deeper nodes exist
    assert tree.match(a + ["u0t2"], 1) == 1
