"""BASELINE config definitions used by tools/configs_bench.py, checked against SURVEY.md §8."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load():
    spec = importlib.util.spec_from_file_location("configs_bench", os.path.join(ROOT, "tools", "configs_bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_gpt2_small_state_dict_shapes_match_survey():
    shapes = _load().gpt2_small_shapes()
    numel = [int(__import__("math").prod(s)) for s in shapes]
    assert len(shapes) == 149                       # 149 tensors incl. the tied lm_head alias
    assert sum(numel) == 163_037_184                # 652 MB in fp32
    assert sum(1 for n in numel if n * 4 <= 12 * 1024) == 98   # "98 of 149 messages are <= 12 KB"
    assert max(numel) * 4 == 154_389_504            # 2 x 154.4 MB (wte and its alias)


def test_llama_style_parameter_count_matches_survey():
    hidden, layers, kv_heads, heads, mlp, vocab = 2048, 16, 8, 32, 8192, 128256
    head_dim = hidden // heads
    per_layer = (hidden * hidden            # q
                 + 2 * hidden * kv_heads * head_dim   # k, v
                 + hidden * hidden          # o
                 + 3 * hidden * mlp         # gate, up, down
                 + 2 * hidden)              # two RMSNorm weights
    total = vocab * hidden + layers * per_layer + hidden   # tied embeddings + final norm
    assert total == 1_235_814_400           # SURVEY §8(d) cfg4; tools/configs_bench.py uses this count
