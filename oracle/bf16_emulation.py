"""Idealised bf16-operand pipeline on the CPU -- TEST INFRASTRUCTURE (oracle), not product code.

Every matrix product of the reference path (oracle/m2fnet_oracle.py `forward`) is re-run with its two operands rounded to
bf16 -- in the forward and in both backward products -- while everything else (accumulation, LayerNorm, softmax, residual
stream, loss) stays in the caller's dtype (fp64 in the fixtures).  This is the best any "bf16 operands, fp32 accumulate"
implementation can do, so its error against the fp64 reference is the yardstick the product's bf16 mode is held to
(tests/test_gpu_model.py): BASELINE.json's 2e-2 is met on logits, but on gradients it is not reachable with plain bf16
operands at MELD batch sizes (SURVEY.md H2 measured the same for torch's own autocast); the 2-plane mode (bf16x2) meets it.
"""
import math

import torch

from oracle import m2fnet_oracle as O

def r(x): return x.to(torch.bfloat16).to(x.dtype)
class RM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w):   # x [..., K], w [N, K]
        xr, wr = r(x), r(w); ctx.save_for_backward(xr, wr); return xr @ wr.T
    @staticmethod
    def backward(ctx, g):
        xr, wr = ctx.saved_tensors; gr = r(g)
        return gr @ wr, gr.reshape(-1, gr.shape[-1]).T @ xr.reshape(-1, xr.shape[-1])
class RB(torch.autograd.Function):  # batched q k^T style matmul with rounding: a [..,m,k] b [..,k,n]
    @staticmethod
    def forward(ctx, a, b):
        ar, br = r(a), r(b); ctx.save_for_backward(ar, br); return ar @ br
    @staticmethod
    def backward(ctx, g):
        ar, br = ctx.saved_tensors; gr = r(g)
        return gr @ br.transpose(-1,-2), ar.transpose(-1,-2) @ gr
mm = RM.apply
def mha(q_in,k_in,v_in,W,b,Wo,bo,H,keymask):
    B,S,E=q_in.shape; hd=E//H
    q=mm(q_in,W[0:E])+b[0:E]; k=mm(k_in,W[E:2*E])+b[E:2*E]; v=mm(v_in,W[2*E:])+b[2*E:]
    q=q.reshape(B,S,H,hd).permute(0,2,1,3); k=k.reshape(B,S,H,hd).permute(0,2,1,3); v=v.reshape(B,S,H,hd).permute(0,2,1,3)
    s=RB.apply(q,k.transpose(-1,-2))/math.sqrt(hd)
    s=s.masked_fill(keymask[:,None,None,:],float("-inf"))
    P=torch.softmax(s,-1)
    o=RB.apply(P,v).permute(0,2,1,3).reshape(B,S,E)
    return mm(o,Wo)+bo
def forward(params,cfg,text,audio,mask):
    g=lambda n: params[O.canonical_name(n)]
    def modality(x,mod,mc):
        m=O.MOD_NAME[mod]
        for t in range(mc.n_transformers):
            y=x
            for l in range(mc.n_encoder_layers):
                P=f"{m}_encoders.{t}.layers.{l}."
                sa=mha(y,y,y,g(P+"self_attn.in_proj_weight"),g(P+"self_attn.in_proj_bias"),g(P+"self_attn.out_proj.weight"),g(P+"self_attn.out_proj.bias"),mc.n_head,mask)
                y=O.layer_norm(y+sa,g(P+"norm1.weight"),g(P+"norm1.bias"))
                h=torch.relu(mm(y,g(P+"linear1.weight"))+g(P+"linear1.bias"))
                f=mm(h,g(P+"linear2.weight"))+g(P+"linear2.bias")
                y=O.layer_norm(y+f,g(P+"norm2.weight"),g(P+"norm2.bias"))
            x=x+O.layer_norm(y,g(f"{m}_encoders.{t}.norm.weight"),g(f"{m}_encoders.{t}.norm.bias"))
        return mm(x,g(f"{m}_proj.weight"))+g(f"{m}_proj.bias")
    a=modality(audio,0,cfg.AUDIO); t_=modality(text,1,cfg.TEXT)
    for i in range(cfg.FAM.n_layers):
        P=f"fusion_layers.{i}."
        x=mha(t_,a,t_,g(P+"multihead_attention.in_proj_weight"),g(P+"multihead_attention.in_proj_bias"),g(P+"multihead_attention.out_proj.weight"),g(P+"multihead_attention.out_proj.bias"),cfg.FAM.n_head,mask)
        u=torch.relu(torch.cat((x,t_),2)); t_=torch.relu(mm(u,g(P+"linear.weight"))+g(P+"linear.bias"))
    z=torch.cat((a,t_),2)
    idx=O.classifier_linear_indices(cfg.CLASSIFIER.n_layers)
    for j,k in enumerate(idx):
        z=mm(z,g(f"output_layer.{k}.weight"))+g(f"output_layer.{k}.bias")
        if j<len(idx)-1: z=torch.relu(z)
    return z


def loss_and_grads(params, cfg, text, audio, mask, target, weight=None):
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items() if O.canonical_name(k) == k}
    logits = forward(leaves, cfg, text, audio, mask)
    loss = O.ce_loss(logits, target, weight)
    names = list(leaves)
    grads = torch.autograd.grad(loss, [leaves[n] for n in names])
    return logits.detach(), loss.detach(), dict(zip(names, grads))
