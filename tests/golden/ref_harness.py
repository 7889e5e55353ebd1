"""Import the *unmodified* reference (``/root/reference``) on a CPU-only box.

TEST INFRASTRUCTURE ONLY.  Used by ``make_golden.py`` (fixture generation) and by the
``test_live_reference_*`` tests, which skip when ``/root/reference`` is absent (it does
not exist on the GPU box).  Nothing in the product imports this file.

The reference cannot be imported as shipped: ``fairseq`` is not
installed, ``from_pretrained`` needs the network, transformers 5 changed the
``BertLayer``/``ViTLayer`` call signature and ``forward`` hard-codes ``.cuda()``.  This
harness supplies exactly the shims listed in SURVEY.md section 8c and nothing else; the
reference's ``mDT/src/modules/*.py``, ``collator.py`` and ``pre_processing.py`` are
loaded verbatim by path and ``algos.pyx`` is compiled with Cython (language_level=2)
into ``oracle/_ref/``.
"""

from __future__ import annotations

import contextlib
import importlib.util
import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

REF_ROOT = os.environ.get("MDT_REFERENCE_ROOT", "/root/reference")
REF_SRC = os.path.join(REF_ROOT, "mDT", "src")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REF_SRC, "modules"))


# --------------------------------------------------------------------------------------
# fairseq shim: only the symbols the hot-path files import
# --------------------------------------------------------------------------------------
def _install_fairseq_shim() -> None:
    if "fairseq" in sys.modules and getattr(sys.modules["fairseq"], "_mdt_shim", False):
        return
    fairseq = types.ModuleType("fairseq")
    fairseq._mdt_shim = True
    utils = types.ModuleType("fairseq.utils")

    def softmax(x, dim: int, onnx_trace: bool = False):
        return F.softmax(x, dim=dim, dtype=torch.float32)

    def gelu(x):
        return F.gelu(x.float()).type_as(x)

    def get_activation_fn(name: str):
        if name == "relu":
            return F.relu
        if name == "gelu":
            return gelu
        if name == "tanh":
            return torch.tanh
        raise RuntimeError(f"--activation-fn {name} not supported")

    def safe_hasattr(obj, k):
        return getattr(obj, k, None) is not None

    utils.softmax = softmax
    utils.get_activation_fn = get_activation_fn
    utils.safe_hasattr = safe_hasattr

    modules = types.ModuleType("fairseq.modules")
    fdrop = types.ModuleType("fairseq.modules.fairseq_dropout")
    qn = types.ModuleType("fairseq.modules.quant_noise")

    class FairseqDropout(nn.Module):
        def __init__(self, p, module_name=None):
            super().__init__()
            self.p = p
            self.module_name = module_name
            self.apply_during_inference = False

        def forward(self, x, inplace: bool = False):
            if self.p > 0 and (self.training or self.apply_during_inference):
                return F.dropout(x, p=self.p, training=True, inplace=inplace)
            return x

    class LayerDropModuleList(nn.ModuleList):
        def __init__(self, p, modules=None):
            super().__init__(modules)
            self.p = p

    def LayerNorm(normalized_shape, eps=1e-5, elementwise_affine=True, export=False):
        return nn.LayerNorm(normalized_shape, eps, elementwise_affine)

    def quant_noise(module, p, block_size):
        assert p <= 0, "quant noise is off on the hot path"
        return module

    fdrop.FairseqDropout = FairseqDropout
    qn.quant_noise = quant_noise
    modules.FairseqDropout = FairseqDropout
    modules.LayerDropModuleList = LayerDropModuleList
    modules.LayerNorm = LayerNorm
    modules.fairseq_dropout = fdrop
    modules.quant_noise = qn
    fairseq.utils = utils
    fairseq.modules = modules
    sys.modules["fairseq"] = fairseq
    sys.modules["fairseq.utils"] = utils
    sys.modules["fairseq.modules"] = modules
    sys.modules["fairseq.modules.fairseq_dropout"] = fdrop
    sys.modules["fairseq.modules.quant_noise"] = qn


def _load_package(name: str, path: str):
    spec = importlib.util.spec_from_file_location(
        name, os.path.join(path, "__init__.py"), submodule_search_locations=[path]
    )
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def _load_file(name: str, path: str):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


_CACHE = {}


def ref_modules():
    """The reference's ``mDT/src/modules`` package, loaded verbatim."""
    if "modules" not in _CACHE:
        _install_fairseq_shim()
        _CACHE["modules"] = _load_package("_mdt_ref_modules", os.path.join(REF_SRC, "modules"))
    return _CACHE["modules"]


def ref_collator():
    if "collator" not in _CACHE:
        _CACHE["collator"] = _load_file("_mdt_ref_collator", os.path.join(REF_SRC, "data", "collator.py"))
    return _CACHE["collator"]


def ref_pre_processing():
    if "pre" not in _CACHE:
        _CACHE["pre"] = _load_file(
            "_mdt_ref_pre_processing", os.path.join(REF_SRC, "data", "pyg_datasets", "pre_processing.py")
        )
    return _CACHE["pre"]


def ref_algos():
    """Compile the reference ``algos.pyx`` into ``oracle/_ref`` and import it."""
    if "algos" not in _CACHE:
        here = os.path.dirname(os.path.abspath(__file__))
        sys.path.insert(0, os.path.join(here, "..", "..", "oracle"))
        import build_ref  # oracle/build_ref.py

        _CACHE["algos"] = build_ref.load_ref_algos(build=True)
    return _CACHE["algos"]


def ref_hop_functions():
    """``get_relative_depth`` / ``spread_downwards`` lifted from the dataset class source.

    ``hateful_discussions.py`` imports torch_geometric at module level, so the two
    methods (pure Python + ``copy``) are extracted with ``ast`` and executed verbatim.
    """
    if "hops" not in _CACHE:
        import ast
        import copy

        path = os.path.join(REF_ROOT, "mDT", "experiments", "hateful_discussions", "datasets", "hateful_discussions.py")
        src = open(path).read()
        tree = ast.parse(src)
        fns = {}
        for node in ast.walk(tree):
            if isinstance(node, ast.FunctionDef) and node.name in ("get_relative_depth", "spread_downwards"):
                fns[node.name] = ast.get_source_segment(src, node)
        ns2 = {"copy": copy}
        import textwrap

        for fn in fns.values():
            exec(textwrap.dedent(fn), ns2)

        class H:
            pass

        H.get_relative_depth = ns2["get_relative_depth"]
        H.spread_downwards = ns2["spread_downwards"]
        _CACHE["hops"] = H()
    return _CACHE["hops"]


# --------------------------------------------------------------------------------------
# transformers adapters
# --------------------------------------------------------------------------------------
def small_hf_configs(bert_layers=12, vit_layers=12, attention_dropout=0.0, activation_dropout=0.0):
    from transformers import BertConfig, ViTConfig

    bc = BertConfig(
        num_labels=2,
        num_hidden_layers=bert_layers,
        hidden_dropout_prob=activation_dropout,
        attention_probs_dropout_prob=attention_dropout,
    )
    vc = ViTConfig(
        num_hidden_layers=vit_layers,
        hidden_dropout_prob=activation_dropout,
        attention_probs_dropout_prob=attention_dropout,
    )
    bc._attn_implementation = "eager"
    vc._attn_implementation = "eager"
    return bc, vc


@contextlib.contextmanager
def reference_runtime(bert_layers=12, vit_layers=12):
    """Patch (and restore) the three things the reference assumes about its runtime."""
    import transformers
    from transformers import BertForSequenceClassification, ViTModel
    from transformers.models.bert.modeling_bert import BertLayer
    from transformers.models.vit.modeling_vit import ViTLayer

    mods = ref_modules()
    enc_mod = sys.modules["_mdt_ref_modules.multigraphormer_graph_encoder"]

    class _AutoModel:
        @staticmethod
        def from_pretrained(name, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0):
            _, vc = small_hf_configs(bert_layers, vit_layers, attention_probs_dropout_prob, hidden_dropout_prob)
            return ViTModel(vc)

    class _AutoSeq:
        @staticmethod
        def from_pretrained(name, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0):
            bc, _ = small_hf_configs(bert_layers, vit_layers, attention_probs_dropout_prob, hidden_dropout_prob)
            return BertForSequenceClassification(bc)

    old_auto, old_seq = enc_mod.AutoModel, enc_mod.AutoModelForSequenceClassification
    old_bert_fwd, old_vit_fwd = BertLayer.forward, ViTLayer.forward
    old_cuda = torch.Tensor.cuda

    def bert_fwd(self, hidden_states, *args, **kw):
        if len(args) == 6 and args[5] is False:  # legacy 4.x call from multi_graphormer_fusion_layer.py:138-148
            return (old_bert_fwd(self, hidden_states, args[0]),)
        return old_bert_fwd(self, hidden_states, *args, **kw)

    def vit_fwd(self, hidden_states, *args, **kw):
        if len(args) == 2 and args[1] is False:  # legacy call, multi_graphormer_fusion_layer.py:94-96
            return (old_vit_fwd(self, hidden_states),)
        return old_vit_fwd(self, hidden_states, *args, **kw)

    enc_mod.AutoModel, enc_mod.AutoModelForSequenceClassification = _AutoModel, _AutoSeq
    BertLayer.forward, ViTLayer.forward = bert_fwd, vit_fwd
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        yield mods
    finally:
        enc_mod.AutoModel, enc_mod.AutoModelForSequenceClassification = old_auto, old_seq
        BertLayer.forward, ViTLayer.forward = old_bert_fwd, old_vit_fwd
        torch.Tensor.cuda = old_cuda


def nested_tree_from_parent(parent):
    """Reddit-JSON-like nested dict (``{'id', 'tree': [...]}``) for ``get_relative_depth``."""
    nodes = [{"id": i, "tree": []} for i in range(len(parent))]
    for i, p in enumerate(parent):
        if p >= 0:
            nodes[int(p)]["tree"].append(nodes[i])
    return nodes


def reference_head_and_loss(encoder, batched_data, positive_weight=1.5, negative_weight=1.0):
    """``GraphormerEncoder.forward`` head (multi_modal_discussion_transformer.py:256-276) and
    ``node_cross_entropy`` (hatespeech_loss.py:95-118) applied to a reference encoder.

    The two fairseq-registered classes need fairseq's registries, so their ~15 lines of
    arithmetic are applied here to the *reference* encoder's outputs, in fp32.
    """
    text, bn, glob = encoder(batched_data)
    stack = [encoder.text_pooler, encoder.text_dropout, encoder.node_classifier]
    for layer in stack:
        text = layer(text)
        bn = layer(bn)
    out_all = (text + bn) / 2
    logits = out_all[batched_data["y_mask"]].float()
    targets = torch.flatten(batched_data["y"]).long()
    w = torch.tensor([negative_weight, positive_weight], dtype=torch.float32)
    loss = F.cross_entropy(logits, targets, reduction="sum", weight=w)
    return out_all, glob, loss
