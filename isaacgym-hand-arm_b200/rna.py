"""Random network adversary (``isaacgymenvs/utils/rna_util.py:37-150``) for the DeXtreme tasks.

Same class surface as the reference (``RandomNetworkAdversary(num_envs, in_dims, out_dims, softmax_bins, device)``, ``_refresh``,
``refresh_dropout_masks``, ``fc1 .. fc3`` weights, ``dropout_masks1/2``) so a task or a test can swap it in, plus
``perturbation(dof_state, object_pose_wrt_wrist, discretised_dofs)``, which is what the task needs every step
(``allegro_hand_dextreme.py:864-879``): the DOF-space value of the most likely bin per joint.  Softmax is monotone, so no
probabilities are materialised; the permute / gather / permute of the reference collapses into the select kernel.

The thin first layer, the bias / ReLU / dropout-mask epilogues and the argmax + gather tail are hand-written kernels
(``csrc/ha_rna.cu``); the four wide layers (512x512, 512x1024, 1024x1024, 1024x512) are fp32 library GEMMs (``torch.mm`` ->
cuBLAS) between them.  No CPU fallback: tensors must live on a CUDA device."""
from __future__ import annotations

import ctypes
import math

import torch

from . import _lib

C = _lib.CONSTS


class _Linear:
    """Weight holder with nn.Linear's attribute names (``weight`` [out, in], ``bias`` [out])."""

    def __init__(self, fan_in: int, fan_out: int, device, gen: torch.Generator):
        bound = 1.0 / math.sqrt(fan_in)                                   # nn.Linear's bias init; the weights are re-drawn by _init_weights
        self.weight = torch.empty(fan_out, fan_in, device=device)
        self.bias = ((torch.rand(fan_out, generator=gen) * 2 - 1) * bound).to(device)


class RandomNetworkAdversary:
    def __init__(self, num_envs: int, in_dims: int, out_dims: int, softmax_bins: int, device, seed: int = 0, impl: str = "tc"):
        """``impl``: "tc" = hand-written tcgen05 layers on split-bf16 operands (``csrc/ha_rna_tc.cu``, the default);
        "fp32" = fp32 library GEMMs between the hand-written ends (the first version, kept as the cross-check)."""
        assert impl in ("tc", "fp32")
        self.impl = impl
        if torch.device(device).type != "cuda":
            raise _lib.HandArmLibraryError("RandomNetworkAdversary needs a CUDA device: there is no CPU fallback")
        assert in_dims == C["HA_RNA_IN_DIMS"] and out_dims == C["HA_RNA_DOFS"] and softmax_bins == C["HA_RNA_BINS"], \
            "the kernels are built for the Allegro hand: 16 joints + cube pose in, 16 x 32 bins out"
        self._lib = _lib.load()
        self.in_dims, self.out_dims, self.softmax_bins, self.num_envs, self.device = in_dims, out_dims, softmax_bins, num_envs, device
        self.num_feats1, self.num_feats2 = C["HA_RNA_FEATS1"], C["HA_RNA_FEATS2"]
        self._gen = torch.Generator().manual_seed(seed)
        g = self._gen
        self.fc1 = _Linear(in_dims, self.num_feats1, device, g)
        self.fc1_1 = _Linear(self.num_feats1, self.num_feats1, device, g)
        self.fc2 = _Linear(self.num_feats1, self.num_feats2, device, g)
        self.fc2_1 = _Linear(self.num_feats2, self.num_feats2, device, g)
        self.fc3 = _Linear(self.num_feats2, out_dims * softmax_bins, device, g)
        N = num_envs
        f = dict(dtype=torch.float32, device=device)
        self.dropout_masks1 = torch.zeros(N, self.num_feats1, **f)
        self.dropout_masks2 = torch.zeros(N, self.num_feats2, **f)
        # activations, reused every step
        self._h1 = torch.empty(N, self.num_feats1, **f)
        self._h1b = torch.empty(N, self.num_feats1, **f)
        self._h2 = torch.empty(N, self.num_feats2, **f)
        self._h2b = torch.empty(N, self.num_feats2, **f)
        self._logits = torch.empty(N, out_dims * softmax_bins, **f)
        bf = dict(dtype=torch.bfloat16, device=device)
        self._px = [torch.empty(N, C["HA_RNA_IN_PAD"], **bf) for _ in range(2)]   # hi / lo planes of the padded input rows
        self._p1 = [torch.empty(N, self.num_feats1, **bf) for _ in range(4)]        # hi / lo planes of the 512-wide activations (two layers)
        self._p2 = [torch.empty(N, self.num_feats2, **bf) for _ in range(4)]        # ... of the 1024-wide activations
        self.out = torch.empty(N, out_dims, **f)
        self.bins = torch.empty(N, out_dims, dtype=torch.int32, device=device)
        self._refresh()

    # ---- rna_util.py:96-122 ----------------------------------------------------------------------------------------------
    def _refresh(self):
        self._init_weights()
        self.refresh_dropout_masks()

    def _init_weights(self):
        """nn.init.kaiming_uniform_(w) with its defaults (a = 0, fan_in, leaky_relu): U(-b, b), b = sqrt(6 / fan_in)."""
        for lin in (self.fc1, self.fc1_1, self.fc2, self.fc2_1, self.fc3):
            bound = math.sqrt(6.0 / lin.weight.shape[1])
            w = (torch.rand(lin.weight.shape, generator=self._gen) * 2 - 1) * bound
            lin.weight.copy_(w)
        self._wt = None
        self._wsplit = None

    def refresh_dropout_masks(self):
        p = torch.rand(2, generator=self._gen)
        m1 = torch.bernoulli(torch.full((self.num_envs, self.num_feats1), float(p[0])), generator=self._gen)
        m2 = torch.bernoulli(torch.full((self.num_envs, self.num_feats2), float(p[1])), generator=self._gen)
        self.dropout_masks1.copy_(m1)
        self.dropout_masks2.copy_(m2)
        self._mbits = None

    def load_from(self, ref) -> None:
        """Copy weights and masks from a reference ``RandomNetworkAdversary`` (any device)."""
        for name in ("fc1", "fc1_1", "fc2", "fc2_1", "fc3"):
            getattr(self, name).weight.copy_(getattr(ref, name).weight.detach())
            getattr(self, name).bias.copy_(getattr(ref, name).bias.detach())
        self.dropout_masks1.copy_(ref.dropout_masks1)
        self.dropout_masks2.copy_(ref.dropout_masks2)
        self._wt = None
        self._wsplit = None
        self._mbits = None

    # ---- forward ----------------------------------------------------------------------------------------------------------
    def _weights_t(self):
        if self._wt is None:                                              # [in, out] row-major operands of the library GEMMs
            self._wt = [l.weight.t().contiguous() for l in (self.fc1_1, self.fc2, self.fc2_1, self.fc3)]
        return self._wt

    def _logits_into(self, dof_state: torch.Tensor, object_pose_wrt_wrist: torch.Tensor) -> torch.Tensor:
        lib, N = self._lib, self.num_envs
        assert dof_state.is_contiguous() and dof_state.numel() == N * 16 * 2
        assert object_pose_wrt_wrist.shape == (N, 7) and object_pose_wrt_wrist.stride(1) == 1, "pose rows must be 7 consecutive floats"
        st = ctypes.c_void_p(torch.cuda.current_stream(self.out.device).cuda_stream)
        P = lambda t: ctypes.c_void_p(t.data_ptr())
        w11, w2, w21, w3 = self._weights_t()
        _lib.check(lib.ha_rna_input_layer(N, P(dof_state), P(object_pose_wrt_wrist), object_pose_wrt_wrist.stride(0), P(self.fc1.weight), P(self.fc1.bias), P(self._h1),
                                          None, None, st))
        torch.mm(self._h1, w11, out=self._h1b)                                                        # fc1_1
        _lib.check(lib.ha_rna_bias_act(P(self._h1b), P(self.fc1_1.bias), P(self.dropout_masks1), 0, N, self.num_feats1, st))
        torch.mm(self._h1b, w2, out=self._h2)                                                         # fc2 + ReLU
        _lib.check(lib.ha_rna_bias_act(P(self._h2), P(self.fc2.bias), None, 1, N, self.num_feats2, st))
        torch.mm(self._h2, w21, out=self._h2b)                                                        # fc2_1
        _lib.check(lib.ha_rna_bias_act(P(self._h2b), P(self.fc2_1.bias), P(self.dropout_masks2), 0, N, self.num_feats2, st))
        torch.mm(self._h2b, w3, out=self._logits)                                                     # fc3 (bias joins in the select kernel)
        return self._logits

    def _weights_split(self):
        """bf16 hi / lo planes of the four wide weight matrices ([out, in] row-major, as nn.Linear stores them)."""
        if self._wsplit is None:
            st = ctypes.c_void_p(torch.cuda.current_stream(self.out.device).cuda_stream)
            self._wsplit = []
            w1p = torch.zeros(self.num_feats1, C["HA_RNA_IN_PAD"], dtype=torch.float32, device=self.fc1.weight.device)
            w1p[:, : self.in_dims] = self.fc1.weight                        # fc1 with K padded to one tensor-core k-block
            for w in (w1p, self.fc1_1.weight, self.fc2.weight, self.fc2_1.weight, self.fc3.weight):
                w = w.contiguous()
                hi, lo = torch.empty_like(w, dtype=torch.bfloat16), torch.empty_like(w, dtype=torch.bfloat16)
                _lib.check(self._lib.ha_rna_split_bf16(ctypes.c_void_p(w.data_ptr()), ctypes.c_void_p(hi.data_ptr()), ctypes.c_void_p(lo.data_ptr()),
                                                       w.numel(), st))
                self._wsplit.append((hi, lo))
        return self._wsplit

    def _mask_bits(self):
        """The {0, 1} dropout masks packed 32 columns to a word (bit j = column 32 c + j): 2 MB instead of 100 MB read per step."""
        if self._mbits is None:
            w = (1 << torch.arange(32, dtype=torch.int64, device=self.dropout_masks1.device))
            self._mbits = []
            for m in (self.dropout_masks1, self.dropout_masks2):
                words = ((m.view(m.shape[0], -1, 32) != 0).to(torch.int64) * w).sum(-1)            # [N, F / 32] in 0 .. 2^32 - 1
                self._mbits.append(torch.where(words >= 2 ** 31, words - 2 ** 32, words).to(torch.int32).contiguous())
        return self._mbits

    def _perturbation_tc(self, dof_state, object_pose_wrt_wrist, discretised_dofs) -> torch.Tensor:
        lib, N = self._lib, self.num_envs
        assert dof_state.is_contiguous() and dof_state.numel() == N * 16 * 2
        assert object_pose_wrt_wrist.shape == (N, 7) and object_pose_wrt_wrist.stride(1) == 1, "pose rows must be 7 consecutive floats"
        st = ctypes.c_void_p(torch.cuda.current_stream(self.out.device).cuda_stream)
        P = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
        (w1h, w1l), (w11h, w11l), (w2h, w2l), (w21h, w21l), (w3h, w3l) = self._weights_split()
        mb1, mb2 = self._mask_bits()
        a_h, a_l, b_h, b_l = self._p1
        c_h, c_l, d_h, d_l = self._p2
        F1, F2 = self.num_feats1, self.num_feats2
        x_h, x_l = self._px
        _lib.check(lib.ha_rna_input_planes(N, P(dof_state), P(object_pose_wrt_wrist), object_pose_wrt_wrist.stride(0), P(x_h), P(x_l), st))
        _lib.check(lib.ha_rna_layer_tc(N, F1, C["HA_RNA_IN_PAD"], P(x_h), P(x_l), P(w1h), P(w1l), P(self.fc1.bias), None, 1, P(a_h), P(a_l), st))   # fc1 + ReLU
        _lib.check(lib.ha_rna_layer_tc(N, F1, F1, P(a_h), P(a_l), P(w11h), P(w11l), P(self.fc1_1.bias), P(mb1), 0,
                                       P(b_h), P(b_l), st))                                                                # fc1_1, * mask1
        _lib.check(lib.ha_rna_layer_tc(N, F2, F1, P(b_h), P(b_l), P(w2h), P(w2l), P(self.fc2.bias), None, 1, P(c_h), P(c_l), st))      # fc2 + ReLU
        _lib.check(lib.ha_rna_layer_tc(N, F2, F2, P(c_h), P(c_l), P(w21h), P(w21l), P(self.fc2_1.bias), P(mb2), 0,
                                       P(d_h), P(d_l), st))                                                                # fc2_1, * mask2
        _lib.check(lib.ha_rna_select_tc(N, F2, P(d_h), P(d_l), P(w3h), P(w3l), P(self.fc3.bias), P(discretised_dofs), P(self.out), P(self.bins), st))
        return self.out

    def perturbation(self, dof_state: torch.Tensor, object_pose_wrt_wrist: torch.Tensor, discretised_dofs: torch.Tensor) -> torch.Tensor:
        """[N,16] joint-space value of the most likely bin per joint (``allegro_hand_dextreme.py:871-876``)."""
        if self.impl == "tc":
            return self._perturbation_tc(dof_state, object_pose_wrt_wrist, discretised_dofs)
        logits = self._logits_into(dof_state, object_pose_wrt_wrist)
        st = ctypes.c_void_p(torch.cuda.current_stream(self.out.device).cuda_stream)
        P = lambda t: ctypes.c_void_p(t.data_ptr())
        _lib.check(self._lib.ha_rna_select(self.num_envs, P(logits), P(self.fc3.bias), P(discretised_dofs), P(self.out), P(self.bins), st))
        return self.out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """Reference surface (``rna_util.py:124-143``): softmax probabilities [N, 16, 32] for a concatenated input [N, 23]."""
        dof_state = torch.stack([x[:, :16], torch.zeros_like(x[:, :16])], dim=-1).contiguous()
        logits = self._logits_into(dof_state, x[:, 16:].contiguous()) + self.fc3.bias
        return torch.softmax(logits.view(-1, self.out_dims, self.softmax_bins), dim=-1)

    __call__ = forward

    def flops_per_env(self) -> int:
        f1, f2, o = self.num_feats1, self.num_feats2, self.out_dims * self.softmax_bins
        return 2 * (self.in_dims * f1 + f1 * f1 + f1 * f2 + f2 * f2 + f2 * o)
