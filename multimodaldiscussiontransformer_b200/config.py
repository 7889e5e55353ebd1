"""Hyper-parameters of the reference's published run as an ``args`` namespace
(mDT/experiments/hateful_discussions/run_train.sh:28-65 called as ``run_train.sh 8 4 5 2 2 0``,
sample_run.sh:3): what ``fairseq-train`` would hand to ``GraphormerModel.build_model``."""

from __future__ import annotations

from argparse import Namespace

PUBLISHED = dict(
    arch="multi_graphormer_base",
    dropout=0.4, attention_dropout=0.3, act_dropout=0.3,
    encoder_layers=4, num_bottleneck_tokens=4, num_fusion_layers=8, num_graph_stack=2, num_fusion_stack=2,
    encoder_embed_dim=768, encoder_ffn_embed_dim=768, encoder_attention_heads=12,
    freeze_initial_encoders=True, num_classes=1, max_nodes=10000,
    # criterion / optimizer / data side
    positive_weight=1.5, negative_weight=1.0, lr=3e-5, end_learning_rate=3e-7, warmup_updates=3246,
    total_num_update=10820, weight_decay=0.01, adam_betas=(0.9, 0.999), adam_eps=1e-8, update_freq=3, batch_size=12,
    spatial_pos_max=5,
)


def published_args(**overrides) -> Namespace:
    d = dict(PUBLISHED)
    d.update(overrides)
    return Namespace(**d)
