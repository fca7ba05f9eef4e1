"""Drop-in mirror of vae_npvc/model/layers_gst.py: ``StyleTokenLayer`` and ``MultiHeadedAttention``.

Same constructor arguments and parameter names (``gst_embs``, ``mha.linear_{q,k,v,out}.{weight,bias}``).  The modules are
parameter holders; forward runs the CUDA path: four small linears (vnb_linear_fwd / bwd), the attention core
(vnb_gst_attn_fwd / bwd, csrc/hier.cu).  tanh of the (tokens x dim) table is a parameter-sized torch op, like the
row-normalisation of a normalised speaker table in layers.Conditions.
"""
import torch
import torch.nn as nn

from .. import functional as Fn


class MultiHeadedAttention(nn.Module):
    """Reference layers_gst.py:63-147 for the one use the model makes of it: one query per utterance (time1 == 1), the
    same keys / values for every utterance, no mask, dropout 0."""

    def __init__(self, q_dim, k_dim, v_dim, n_head, n_feat, dropout_rate=0.0):
        super(MultiHeadedAttention, self).__init__()
        assert n_feat % n_head == 0
        assert dropout_rate == 0.0, 'dropout is 0.0 in the shipped recipe; not on the B200 hot path'
        self.d_k = n_feat // n_head
        self.h = n_head
        self.linear_q = nn.Linear(q_dim, n_feat)
        self.linear_k = nn.Linear(k_dim, n_feat)
        self.linear_v = nn.Linear(v_dim, n_feat)
        self.linear_out = nn.Linear(n_feat, n_feat)
        self.attn = None
        self.dropout = nn.Dropout(p=dropout_rate)

    def forward(self, query, key, value, mask=None):
        """query (B, 1, q_dim) or (B, q_dim); key / value (N, k_dim) shared by all utterances (an expanded
        (B, N, k_dim) is accepted and its first entry used).  -> (B, 1, n_feat) / (B, n_feat)"""
        assert mask is None, 'masks are not used by StyleTokenLayer'
        squeeze = query.dim() == 3
        q_in = query[:, 0] if squeeze else query
        if key.dim() == 3:
            key, value = key[0], value[0]
        q = Fn.LinearFn.apply(q_in, self.linear_q.weight, self.linear_q.bias)
        k = Fn.LinearFn.apply(key, self.linear_k.weight, self.linear_k.bias)
        v = Fn.LinearFn.apply(value, self.linear_v.weight, self.linear_v.bias)
        ctx = Fn.GstAttnFn.apply(q, k, v, self.h)
        out = Fn.LinearFn.apply(ctx, self.linear_out.weight, self.linear_out.bias)
        return out.unsqueeze(1) if squeeze else out


class StyleTokenLayer(nn.Module):
    """Reference layers_gst.py:10-60."""

    def __init__(self, ref_embed_dim=128, gst_tokens=10, gst_token_dim=256, gst_heads=4, dropout_rate=0.0):
        super(StyleTokenLayer, self).__init__()
        gst_embs = torch.randn(gst_tokens, gst_token_dim // gst_heads)
        self.register_parameter('gst_embs', nn.Parameter(gst_embs))
        self.mha = MultiHeadedAttention(q_dim=ref_embed_dim, k_dim=gst_token_dim // gst_heads,
                                        v_dim=gst_token_dim // gst_heads, n_head=gst_heads, n_feat=gst_token_dim,
                                        dropout_rate=dropout_rate)

    def forward(self, ref_embs):
        """ref_embs (B, ref_embed_dim) -> style embedding (B, gst_token_dim)."""
        tok = torch.tanh(self.gst_embs)
        return self.mha(ref_embs, tok, tok, None)
