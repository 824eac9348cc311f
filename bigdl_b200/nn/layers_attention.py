"""Attention, FeedForwardNetwork, Transformer (language model + translation) and SequenceBeamSearch.

Reference: S/nn/Attention.scala:30-160 (bias-free Q/K/V/output dense layers, split heads, scaled QK^T + bias, softmax,
dropout, PV, combine heads; K/V cache for incremental decoding), FeedForwardNetwork.scala, Transformer.scala:53-420
(tensor2tensor-style pre-norm blocks: x + dropout(sublayer(layer_norm(x))), sinusoid position signal, optional output
projection sharing the embedding matrix), TransformerOperation.scala, SequenceBeamSearch.scala.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from ..utils.table import Table, to_table
from .layers_core import Dropout, Linear
from .layers_linear import LayerNormalization, LookupTable
from .module import Container, Module


def position_signal(length: int, channels: int, device, dtype=torch.float32, min_timescale=1.0, max_timescale=1.0e4,
                    start: int = 0) -> torch.Tensor:
    """Sinusoid timing signal [length, channels] (TransformerOperation.addTimingSignal1D)."""
    pos = torch.arange(start, start + length, device=device, dtype=torch.float32)
    num = channels // 2
    log_inc = math.log(max_timescale / min_timescale) / max(num - 1, 1)
    inv = min_timescale * torch.exp(torch.arange(num, device=device, dtype=torch.float32) * -log_inc)
    scaled = pos[:, None] * inv[None, :]
    sig = torch.cat([torch.sin(scaled), torch.cos(scaled)], dim=1)
    if channels % 2:
        sig = F.pad(sig, (0, 1))
    return sig.to(dtype)


def causal_bias(length: int, device, dtype=torch.float32) -> torch.Tensor:
    """[1, 1, L, L] additive mask with -1e9 above the diagonal (TransformerOperation.attentionBiasLowerTriangle)."""
    m = torch.full((length, length), -1e9, device=device, dtype=torch.float32).triu(1)
    m = m.reshape(1, 1, length, length).to(dtype)
    m._bdl_causal = True        # lets the fused kernel mask by index (and skip whole tiles) instead of reading L*L floats
    return m


def padding_bias(ids: torch.Tensor, padding_value: float, dtype=torch.float32) -> torch.Tensor:
    """[B, 1, 1, L] additive mask with -1e9 at padded positions (TransformerOperation.getPaddingBias)."""
    return ((ids == padding_value).to(torch.float32) * -1e9).reshape(ids.shape[0], 1, 1, ids.shape[1]).to(dtype)


class Attention(Container):
    """Multi-head attention. input = table(x [B,Lq,H], y [B,Lk,H], bias broadcastable to [B,heads,Lq,Lk]); returns
    [B,Lq,H]. ``set_cache``/``clear_cache`` keep K/V across incremental decoding steps (Attention.scala:117-160)."""

    def __init__(self, hidden_size: int, num_heads: int, attention_dropout: float):
        super().__init__()
        if hidden_size % num_heads:
            raise ValueError("hidden size must be evenly divisible by the number of heads")
        self.hidden_size, self.num_heads, self.attention_dropout = hidden_size, num_heads, attention_dropout
        self.q = Linear(hidden_size, hidden_size, False).setName("q")
        self.k = Linear(hidden_size, hidden_size, False).setName("k")
        self.v = Linear(hidden_size, hidden_size, False).setName("v")
        self.o = Linear(hidden_size, hidden_size, False).setName("output_transform")
        self.drop = Dropout(attention_dropout)
        self.modules = [self.q, self.k, self.v, self.o, self.drop]
        self._cache = None

    def set_cache(self, cache: dict | None):
        self._cache = cache
        return self

    def _split(self, x):
        B, L, _ = x.shape
        return x.reshape(B, L, self.num_heads, self.hidden_size // self.num_heads).transpose(1, 2)

    def updateOutput(self, input):
        xs = to_table(input).to_list()
        x, y = xs[0], xs[1]
        bias = xs[2] if len(xs) > 2 else None
        qp, kp, vp = self.q._call(x), self.k._call(y), self.v._call(y)
        if self._cache is None and not (self.train and self.attention_dropout > 0) and qp.dim() == 3:
            from ..ops import attention as fused
            if fused.supported(qp, kp, vp, self.num_heads) and (bias is None or not bias.requires_grad):
                # one tcgen05 kernel: QK^T, bias, softmax and PV with the logits kept on chip (csrc/attention.cu)
                causal = bool(getattr(bias, "_bdl_causal", False)) and qp.shape[1] == kp.shape[1]
                return self.o._call(fused.attention(qp, kp, vp, None if causal else bias, self.num_heads, causal=causal))
        q, k, v = self._split(qp), self._split(kp), self._split(vp)
        if self._cache is not None:
            if "k" in self._cache:
                k = torch.cat([self._cache["k"], k], dim=2)
                v = torch.cat([self._cache["v"], v], dim=2)
            self._cache["k"], self._cache["v"] = k.detach(), v.detach()
        depth = self.hidden_size // self.num_heads
        logits = torch.matmul(q * depth ** -0.5, k.transpose(-1, -2))
        if bias is not None:
            logits = logits + bias.to(logits.dtype)
        w = F.softmax(logits.float(), dim=-1).to(q.dtype)
        w = self.drop._call(w)
        out = torch.matmul(w, v).transpose(1, 2).reshape(x.shape[0], x.shape[1], self.hidden_size)
        return self.o._call(out)


class FeedForwardNetwork(Container):
    """dense(filter, relu) → dropout → dense(hidden) (S/nn/FeedForwardNetwork.scala)."""

    def __init__(self, hidden_size: int, filter_size: int, relu_dropout: float):
        super().__init__()
        self.hidden_size, self.filter_size = hidden_size, filter_size
        self.filter_layer = Linear(hidden_size, filter_size).setName("filter_layer")
        self.output_layer = Linear(filter_size, hidden_size).setName("output_layer")
        self.drop = Dropout(relu_dropout)
        self.modules = [self.filter_layer, self.drop, self.output_layer]

    def updateOutput(self, input):
        return self.output_layer._call(self.drop._call(F.relu(self.filter_layer._call(input))))


class _Block(Container):
    def __init__(self, hidden, heads, filt, attn_drop, ffn_drop, post_drop, cross: bool):
        super().__init__()
        self.ln1, self.self_attn = LayerNormalization(hidden), Attention(hidden, heads, attn_drop)
        self.cross = cross
        if cross:
            self.ln_c, self.cross_attn = LayerNormalization(hidden), Attention(hidden, heads, attn_drop)
        self.ln2, self.ffn = LayerNormalization(hidden), FeedForwardNetwork(hidden, filt, ffn_drop)
        self.post = Dropout(post_drop)
        self.modules = [self.ln1, self.self_attn, self.ln2, self.ffn, self.post] + \
            ([self.ln_c, self.cross_attn] if cross else [])

    def run(self, x, self_bias, enc=None, enc_bias=None):
        n = self.ln1._call(x)
        x = x + self.post._call(self.self_attn._call(Table(n, n, self_bias)))
        if self.cross:
            n = self.ln_c._call(x)
            x = x + self.post._call(self.cross_attn._call(Table(n, enc, enc_bias)))
        n = self.ln2._call(x)
        return x + self.post._call(self.ffn._call(n))

    def updateOutput(self, input):
        xs = to_table(input).to_list()
        return self.run(*xs)


class TransformerType:
    LanguageModel = "LanguageModel"
    Translation = "Translation"


class Transformer(Container):
    """Transformer LM (input ids [B,L] → hidden states or vocab logits) or translation model (table(src ids, tgt ids))
    (S/nn/Transformer.scala:53-420). Ids are token indices with ``padding_value`` marking padding."""

    def __init__(self, vocab_size, hidden_size, num_heads, filter_size, num_hidden_layers, embedding_dropout,
                 attention_dropout, ffn_dropout, padding_value: float = 0, with_share_weights_linear: bool = False,
                 transformer_type: str = TransformerType.LanguageModel, beam_search=None):
        super().__init__()
        self.vocab_size, self.hidden_size, self.num_heads = vocab_size, hidden_size, num_heads
        self.filter_size, self.num_hidden_layers = filter_size, num_hidden_layers
        self.padding_value, self.with_share_weights_linear = padding_value, with_share_weights_linear
        self.transformer_type, self.beam_search = transformer_type, beam_search
        self.embedding = LookupTable(vocab_size, hidden_size, padding_value=padding_value, mask_zero=True)
        with torch.no_grad():
            self.embedding.weight.normal_(0.0, hidden_size ** -0.5)
        self.emb_drop = Dropout(embedding_dropout)
        translation = transformer_type == TransformerType.Translation
        self.decoder = [_Block(hidden_size, num_heads, filter_size, attention_dropout, ffn_dropout, embedding_dropout,
                               cross=translation) for _ in range(num_hidden_layers)]
        self.encoder = [_Block(hidden_size, num_heads, filter_size, attention_dropout, ffn_dropout, embedding_dropout,
                               cross=False) for _ in range(num_hidden_layers)] if translation else []
        self.dec_norm = LayerNormalization(hidden_size)
        self.enc_norm = LayerNormalization(hidden_size) if translation else None
        self.modules = [self.embedding, self.emb_drop] + self.encoder + self.decoder + [self.dec_norm] + \
            ([self.enc_norm] if translation else [])

    def _embed(self, ids, start: int = 0):
        # the reference embeds 1-based ids with 0 as padding (maskZero); scale by sqrt(hidden) and add the timing signal
        e = self.embedding._call(ids) * (self.hidden_size ** 0.5)
        return e + position_signal(ids.shape[1], self.hidden_size, e.device, e.dtype, start=start)

    def _shift_right(self, x):
        return F.pad(x, (0, 0, 1, 0))[:, :-1]

    def _logits(self, h):
        if not self.with_share_weights_linear:
            return h
        return F.linear(h, self.embedding.weight.to(h.dtype))

    def encode(self, src_ids):
        bias = padding_bias(src_ids, self.padding_value)
        x = self.emb_drop._call(self._embed(src_ids))
        for blk in self.encoder:
            x = blk.run(x, bias)
        return self.enc_norm._call(x), bias

    def decode(self, tgt_emb, enc=None, enc_bias=None):
        L = tgt_emb.shape[1]
        bias = causal_bias(L, tgt_emb.device)
        x = self.emb_drop._call(tgt_emb)
        for blk in self.decoder:
            x = blk.run(x, bias, enc, enc_bias)
        return self.dec_norm._call(x)

    def updateOutput(self, input):
        if self.transformer_type == TransformerType.LanguageModel:
            ids = input
            x = self._shift_right(self._embed(ids))
            return self._logits(self.decode(x))
        xs = to_table(input).to_list()
        src = xs[0]
        enc, enc_bias = self.encode(src)
        if len(xs) > 1 and xs[1] is not None:
            tgt = xs[1]
            x = self._shift_right(self._embed(tgt))
            return self._logits(self.decode(x, enc, enc_bias))
        if self.beam_search is None:
            raise ValueError("translation inference needs a SequenceBeamSearch (beamSearch=...)")
        return self._beam_decode(enc, enc_bias)

    # ---- incremental decoding used by SequenceBeamSearch
    def _symbols_to_logits(self, enc, enc_bias):
        caches = [dict() for _ in self.decoder]

        def fn(ids, step, state):
            # ids [B*beam, step+1] (1-based tokens, first column is the start symbol 0 → zero embedding)
            last = ids[:, -1:]
            e = self.embedding._call(last) * (self.hidden_size ** 0.5)
            e = e + position_signal(1, self.hidden_size, e.device, e.dtype, start=step)
            if step == 0:
                e = torch.zeros_like(e)
            x = e
            for blk, cache in zip(self.decoder, state["caches"]):
                blk.self_attn.set_cache(cache)
                x = blk.run(x, None, state["enc"], state["enc_bias"])
                blk.self_attn.set_cache(None)
            h = self.dec_norm._call(x)
            return F.linear(h[:, 0], self.embedding.weight.to(h.dtype)), state

        return fn, {"caches": caches, "enc": enc, "enc_bias": enc_bias}

    def _beam_decode(self, enc, enc_bias):
        fn, state = self._symbols_to_logits(enc, enc_bias)
        return self.beam_search.search(fn, state, enc.shape[0], enc.device)


class SequenceBeamSearch(Module):
    """Length-normalised beam search (S/nn/SequenceBeamSearch.scala): keeps ``beam_size`` alive and finished
    hypotheses per batch item, score = log_prob / ((5 + len) / 6)^alpha, stops at ``max_decode_length`` or when the
    best alive hypothesis cannot beat the worst finished one. Returns table(sequences [B, beam, L], scores [B, beam])."""

    def __init__(self, vocab_size, beam_size, alpha, max_decode_length, eos_id, padding_value, num_hidden_layers,
                 hidden_size):
        super().__init__()
        self.vocab_size, self.beam_size, self.alpha = vocab_size, beam_size, alpha
        self.max_decode_length, self.eos_id, self.padding_value = max_decode_length, eos_id, padding_value
        self.num_hidden_layers, self.hidden_size = num_hidden_layers, hidden_size
        self._logit_fn = None

    def setLogitFn(self, fn):
        self._logit_fn = fn
        return self

    def _length_norm(self, length):
        return ((5.0 + length) / 6.0) ** self.alpha

    @staticmethod
    def _expand(t, beam):
        return t.unsqueeze(1).expand(t.shape[0], beam, *t.shape[1:]).reshape(t.shape[0] * beam, *t.shape[1:])

    def _reorder_state(self, state, idx):
        def rec(v):
            if isinstance(v, torch.Tensor):
                return v.index_select(0, idx)
            if isinstance(v, dict):
                return {k: rec(x) for k, x in v.items()}
            if isinstance(v, list):
                return [rec(x) for x in v]
            return v
        return rec(state)

    def search(self, logit_fn, state, batch: int, device):
        beam, V, NEG = self.beam_size, self.vocab_size, -1e7
        state = {k: (self._expand(v, beam) if isinstance(v, torch.Tensor) else v) for k, v in state.items()}
        alive_seq = torch.zeros(batch, beam, 1, dtype=torch.long, device=device)
        alive_lp = torch.tensor([0.0] + [NEG] * (beam - 1), device=device).repeat(batch, 1)
        fin_seq = torch.zeros(batch, beam, 1, dtype=torch.long, device=device)
        fin_scores = torch.full((batch, beam), NEG, device=device)
        fin_flags = torch.zeros(batch, beam, dtype=torch.bool, device=device)
        with torch.no_grad():
            for step in range(self.max_decode_length):
                flat_ids = alive_seq.reshape(batch * beam, -1)
                logits, state = logit_fn(flat_ids, step, state)
                lp = F.log_softmax(logits.float(), dim=-1).reshape(batch, beam, V) + alive_lp.unsqueeze(2)
                top_lp, top_idx = lp.reshape(batch, beam * V).topk(2 * beam, dim=1)
                beam_idx, tok = top_idx // V, top_idx % V + 1        # tokens are 1-based
                seqs = torch.cat([alive_seq.gather(1, beam_idx.unsqueeze(2).expand(-1, -1, alive_seq.shape[2])),
                                  tok.unsqueeze(2)], dim=2)
                is_eos = tok == int(self.eos_id)
                # new alive = best non-finished candidates
                a_lp = top_lp + is_eos.float() * NEG
                a_top, a_sel = a_lp.topk(beam, dim=1)
                alive_seq = seqs.gather(1, a_sel.unsqueeze(2).expand(-1, -1, seqs.shape[2]))
                alive_lp = a_top
                src_beam = beam_idx.gather(1, a_sel)
                flat_src = (src_beam + torch.arange(batch, device=device).unsqueeze(1) * beam).reshape(-1)
                state = self._reorder_state(state, flat_src)
                # finished pool
                cand_scores = top_lp / self._length_norm(step + 1) + (~is_eos).float() * NEG
                fin_seq = torch.cat([fin_seq, torch.zeros(batch, beam, 1, dtype=torch.long, device=device)], dim=2)
                all_seq = torch.cat([fin_seq, seqs], dim=1)
                all_scores = torch.cat([fin_scores, cand_scores], dim=1)
                all_flags = torch.cat([fin_flags, is_eos], dim=1)
                f_top, f_sel = all_scores.topk(beam, dim=1)
                fin_seq = all_seq.gather(1, f_sel.unsqueeze(2).expand(-1, -1, all_seq.shape[2]))
                fin_scores, fin_flags = f_top, all_flags.gather(1, f_sel)
                # early stop: best possible alive score cannot beat the worst finished score
                best_alive = alive_lp[:, 0] / self._length_norm(self.max_decode_length)
                worst_fin = torch.where(fin_flags.all(dim=1), fin_scores.min(dim=1).values,
                                        torch.full_like(best_alive, NEG))
                if bool((worst_fin > best_alive).all()):
                    break
        any_fin = fin_flags.any(dim=1)
        L = max(fin_seq.shape[2], alive_seq.shape[2])
        pad = lambda s: F.pad(s, (0, L - s.shape[2]))
        out_seq = torch.where(any_fin.reshape(-1, 1, 1), pad(fin_seq), pad(alive_seq))
        out_scores = torch.where(any_fin.reshape(-1, 1), fin_scores, alive_lp)
        return Table(out_seq, out_scores)

    def updateOutput(self, input):
        if self._logit_fn is None:
            raise ValueError("SequenceBeamSearch needs setLogitFn(fn) or to be driven by a Transformer")
        enc, bias = to_table(input).to_list()[:2]
        return self.search(self._logit_fn, {"enc": enc, "enc_bias": bias}, enc.shape[0], enc.device)


class TransformerOperation:
    """Helper functions of the Transformer (S/nn/TransformerOperation.scala:30-174) under their reference names."""

    @staticmethod
    def getPositionEncode(length: int, hidden_size: int, min_timescale: float = 1.0, max_timescale: float = 1.0e4,
                          device="cpu"):
        return position_signal(length, hidden_size, device, min_timescale=min_timescale, max_timescale=max_timescale)

    @staticmethod
    def attentionBiasLowerTriangle(length: int, device="cpu"):
        return causal_bias(length, device)

    @staticmethod
    def getPaddingBias(ids: torch.Tensor, padding_value: float = 0.0):
        return padding_bias(ids, padding_value)

    @staticmethod
    def shiftRight3D(x: torch.Tensor) -> torch.Tensor:
        """Shift the time axis right by one, zero-filling the first step (decoder input preparation)."""
        out = torch.zeros_like(x)
        out[:, 1:] = x[:, :-1]
        return out

    @staticmethod
    def initRangeTensor(length: int, device="cpu") -> torch.Tensor:
        return torch.arange(length, dtype=torch.float32, device=device)

    @staticmethod
    def dense(input_size: int, output_size: int, bias: bool = True, activation=None, name: str | None = None):
        """Dense layer applied to the last axis of a 3-D tensor (TimeDistributed Linear + optional activation)."""
        from .layers_core import Linear
        from .layers_recurrent import TimeDistributed
        from .module import Sequential

        lin = Linear(input_size, output_size, with_bias=bias)
        if name:
            lin.setName(name)
        seq = Sequential().add(TimeDistributed(lin))
        if activation is not None:
            seq.add(activation)
        return seq


# ------------------------------------------------------------------------------------------------ building blocks
# The reference assembles its Transformer graph from these small modules (Transformer.scala:600-750,
# Attention.scala:217-290, LayerNormalization.scala:60-170). The fused path above does their work inside the attention
# kernel; they exist as layers so graphs written against the reference pieces — and their unit specs — carry over.
class PositionEncode(Module):
    """[B, L, C] → the sinusoid position table [L, C] (no gradient to the input)."""

    def updateOutput(self, input):
        return position_signal(input.shape[1], input.shape[2], input.device, input.dtype)


class PositionEncodeWithShift(Module):
    """Shift the sequence right by one step (zero in front, last step dropped) and add the position table."""

    def updateOutput(self, input):
        shifted = torch.cat([torch.zeros_like(input[:, :1]), input[:, :-1]], 1)
        return shifted + position_signal(input.shape[1], input.shape[2], input.device, input.dtype)


class PaddingMask(Module):
    """Token ids [B, L] → additive bias [B, 1, 1, L] with −1e9 where the id equals ``padding_value`` (0)."""

    def __init__(self, padding_value: float = 0.0):
        super().__init__()
        self.padding_value = padding_value

    def updateOutput(self, input):
        return padding_bias(input.detach(), self.padding_value, torch.float32)


class SelfAttentionMask(Module):
    """[B, L, …] → the [1, 1, L, L] bias that hides future positions."""

    def updateOutput(self, input):
        return causal_bias(input.shape[1], input.device, torch.float32)


class SplitHeads(Module):
    """[B, L, H] → [B, heads, L, H / heads]; ``mul`` scales by (H / heads)^−0.5 (the query scaling)."""

    def __init__(self, hiddenSize: int, numHeads: int, mul: bool = False):
        super().__init__()
        self.hiddenSize, self.numHeads, self.mul = hiddenSize, numHeads, mul

    def updateOutput(self, input):
        b, l, _ = input.shape
        depth = self.hiddenSize // self.numHeads
        out = input.reshape(b, l, self.numHeads, depth).transpose(1, 2)
        return out * depth ** -0.5 if self.mul else out


class CombineHeads(Module):
    """[B, heads, L, D] → [B, L, heads · D]."""

    def updateOutput(self, input):
        b, h, l, d = input.shape
        return input.transpose(1, 2).reshape(b, l, h * d)


class SplitTensor(Module):
    """Tensor → Table of ``num`` equal slices along the (1-based) ``dimension``."""

    def __init__(self, dimension: int, num: int):
        super().__init__()
        self.dimension, self.num = dimension, num

    def updateOutput(self, input):
        d = self.dimension - 1
        return Table(*torch.split(input, input.shape[d] // self.num, d))


class VectorProduct(Module):
    """y = x ⊙ weight + bias over the last dimension, weight = 1 and bias = 0 at start: the affine half of the
    reference's LayerNormalization."""

    def __init__(self, hiddenSize: int):
        super().__init__()
        self.hiddenSize = hiddenSize
        self.register_parameter("weight", torch.ones(hiddenSize))
        self.register_parameter("bias", torch.zeros(hiddenSize))

    def reset(self):
        self.weight.data.fill_(1.0)
        self.bias.data.zero_()
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        return input * self.weight.to(input.dtype) + self.bias.to(input.dtype)


# the two TransformerType case objects by their own names (TransformerOperation.scala:172-176)
LanguageModel = TransformerType.LanguageModel
Translation = TransformerType.Translation
