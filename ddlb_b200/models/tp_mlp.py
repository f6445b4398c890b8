"""A tensor-parallel MLP block built from the two primitives.

This is the composition the primitives exist for (Megatron-style TP with sequence
parallelism; BASELINE.json config #5 names the Llama-3-70B shapes): activations arrive
sequence-sharded ``[m/d, h]``;

    up:    H = act( allgather_rows(X) @ W_up[h, f/d] )      -> tp_columnwise (AG -> GEMM)
    down:  Y = reduce_scatter_rows( H @ W_down[f/d, h] )    -> tp_rowwise   (GEMM -> RS)

and leave sequence-sharded ``[m/d, h]``. Both GEMMs run on the fused sm_100a kernels; the
activation (``relu`` / ``silu`` / ``gelu_tanh``) is applied to the fp32 accumulator inside the
up kernel's epilogue, whose output buffer IS the down kernel's A operand: a forward pass is
exactly two kernel launches, no elementwise kernel and no copy. The
reference has no model code; its closest relatives are the two benchmark primitives
themselves (``ddlb/primitives/TPColumnwise``, ``ddlb/primitives/TPRowwise``).
"""
from __future__ import annotations

from typing import Optional

LLAMA3_70B_MLP = {"hidden": 8192, "ffn": 28672}


class TPMLP:
    """``y_shard = down(act(up(allgather(x_shard))))`` on ``world`` ranks.

    ``w_up`` is this rank's ``[hidden, ffn/world]`` column shard, ``w_down`` its
    ``[ffn/world, hidden]`` row shard (both row-major). ``heap`` is the rank's symmetric
    heap (``None`` for one rank).
    """

    def __init__(self, heap, rank: int, world: int, tokens: int, hidden: int, ffn: int, dtype,
                 w_up, w_down, activation: Optional[str] = "silu", **kernel_opts):
        import torch
        from ..ops.tp_gemm import FusedAllGatherGemm, FusedGemmReduceScatter
        if ffn % world or tokens % world:
            raise ValueError("ffn and tokens must be divisible by world")
        f_local = ffn // world
        if tuple(w_up.shape) != (hidden, f_local) or tuple(w_down.shape) != (f_local, hidden):
            raise ValueError(
                f"expected w_up {(hidden, f_local)} and w_down {(f_local, hidden)}, got "
                f"{tuple(w_up.shape)} and {tuple(w_down.shape)}")
        self.rank, self.world = rank, world
        self.tokens, self.hidden, self.ffn = tokens, hidden, ffn
        self.activation = activation
        # H[tokens, f_local]: written by the up kernel's epilogue, read by the down kernel's TMA
        self._h = torch.empty(tokens, f_local, dtype=dtype, device=w_up.device)
        # up projection: global problem [tokens, hidden] x [hidden, f_local]
        self.up = FusedAllGatherGemm(heap, rank, world, tokens, f_local, hidden, dtype, w_up,
                                     order="AG_before", activation=activation, out=self._h,
                                     **kernel_opts)
        # down projection: every rank holds H[tokens, f_local]; k of the global problem = ffn
        self.down = FusedGemmReduceScatter(heap, rank, world, tokens, hidden, ffn, dtype,
                                           self._h, w_down, **kernel_opts)
        self.x_shard = self.up.a_shard             # write the rank's activations here

    def forward(self, x_shard=None):
        """Returns this rank's ``[tokens/world, hidden]`` rows of the block output.

        ``x_shard`` may be overwritten as soon as the previous ``forward`` has completed on
        this rank's stream: the up kernel ends only after every peer has finished reading it.
        """
        if x_shard is not None:
            self.x_shard.copy_(x_shard)
        self.up.run()                              # H = act(allgather(X) @ W_up), one kernel
        return self.down.run()                     # [tokens/world, hidden], one kernel

    __call__ = forward

    @property
    def gpu_launches(self) -> int:
        return self.up.launches + self.down.launches

    def flops(self) -> float:
        return 2.0 * 2.0 * self.tokens * self.hidden * self.ffn

    def close(self) -> None:
        self.up.close()
        self.down.close()
