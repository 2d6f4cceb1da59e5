"""diffusers 0.27.2 Attention with the default AttnProcessor2_0, self-attention case only:
to_q/to_k/to_v Linear(bias=bias), to_out=[Linear(inner,query_dim,bias=True), Dropout],
scale=dim_head**-0.5, float attention masks are ADDED to the scores (SDPA semantics)."""
import torch
import torch.nn as nn
import torch.nn.functional as F
class Attention(nn.Module):
    def __init__(self, query_dim, cross_attention_dim=None, heads=8, dim_head=64, dropout=0.0,
                 bias=False, upcast_attention=False, **kw):
        super().__init__()
        inner = heads * dim_head
        self.heads = heads
        self.scale = dim_head ** -0.5
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(cross_attention_dim or query_dim, inner, bias=bias)
        self.to_v = nn.Linear(cross_attention_dim or query_dim, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim, bias=True), nn.Dropout(dropout)])
    def forward(self, hidden_states, encoder_hidden_states=None, attention_mask=None, **kw):
        B, S, _ = hidden_states.shape
        ctx = hidden_states if encoder_hidden_states is None else encoder_hidden_states
        q = self.to_q(hidden_states); k = self.to_k(ctx); v = self.to_v(ctx)
        H = self.heads; D = q.shape[-1] // H
        q = q.view(B, -1, H, D).transpose(1, 2)
        k = k.view(B, -1, H, D).transpose(1, 2)
        v = v.view(B, -1, H, D).transpose(1, 2)
        if attention_mask is not None:
            # prepare_attention_mask: repeat_interleave(heads, dim=0) then view(B, heads, -1, S_k)
            attention_mask = attention_mask.repeat_interleave(H, dim=0).view(B, H, -1, attention_mask.shape[-1])
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=attention_mask, dropout_p=0.0, is_causal=False)
        o = o.transpose(1, 2).reshape(B, -1, H * D).to(q.dtype)
        o = self.to_out[0](o)
        o = self.to_out[1](o)
        return o
