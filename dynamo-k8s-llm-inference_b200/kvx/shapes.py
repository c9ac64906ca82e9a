"""KV-cache shapes of the models the reference's manifests and scripts name.

meta-llama/Llama-3.2-1B-Instruct: examples/deploy/{vllm,sglang,trtllm}/disagg.yaml (args --model);
Qwen/Qwen3-0.6B: chat.sh:4, examples/dgdr/trtllm/disagg.yaml:36.  The 8B / 70B shapes belong to
BASELINE.json configs #3-#5.  Values are the public Hugging Face config fields (num_hidden_layers,
num_key_value_heads, head_dim); they are not in the reference and are printed in
every result file.
"""
from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class KVShape:
    name: str
    layers: int
    kv_heads: int
    head_dim: int
    elem_bytes: int = 2  # bf16
    block_tokens: int = 16  # examples/deploy/sglang/disagg.yaml:39-40 (--page-size 16)

    @property
    def regions(self) -> int:
        return 2 * self.layers

    def chunk_bytes(self, tp_size: int = 1) -> int:
        return self.block_tokens * max(1, self.kv_heads // tp_size) * self.head_dim * self.elem_bytes

    def blocks_for(self, isl: int) -> int:
        return -(-isl // self.block_tokens)


KV_SHAPES = {
    "meta-llama/Llama-3.2-1B-Instruct": KVShape("meta-llama/Llama-3.2-1B-Instruct", 16, 8, 64),
    "Qwen/Qwen3-0.6B": KVShape("Qwen/Qwen3-0.6B", 28, 8, 128),
    "meta-llama/Llama-3.1-8B": KVShape("meta-llama/Llama-3.1-8B", 32, 8, 128),
    "meta-llama/Llama-3.1-70B": KVShape("meta-llama/Llama-3.1-70B", 80, 8, 128),
}
KV_SHAPES["meta-llama/Llama-3.1-8B-Instruct"] = KV_SHAPES["meta-llama/Llama-3.1-8B"]
KV_SHAPES["meta-llama/Llama-3.1-70B-Instruct"] = KV_SHAPES["meta-llama/Llama-3.1-70B"]


def payload_bytes(shape: KVShape, isl: int, tp_size: int = 1) -> int:
    """B = n * 2L * chunk: bytes one rank holds for one request."""
    return shape.blocks_for(isl) * shape.regions * shape.chunk_bytes(tp_size)
