"""Full-width decoder layers at the benchmark's shape (B=8 clips x S=2048 tokens -> M=16384 rows, hidden 4096, MLP 14336,
GQA 32:8, LoRA r=16 on all 7 projections) against Hugging Face `MistralModel` + the restated peft `lora.Linear` in fp32 on
the same GPU -- the CTA-pair GEMMs with the fused LoRA k-block, the SwiGLU epilogues, causal flash attention forward /
backward and the rank-r gradient GEMMs exactly as bench.py launches them. Two layers, so that layer 1 also runs the dX of
the q/k/v projection and the first RMSNorm backward (layer 0's input gradient is skipped: the embeddings are frozen).
Reference: videollama2_mistral.py:63-108 -> MistralForCausalLM.forward, litmodule.py:101-120 (LoRA wrap)."""
import types

import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().float(), b.detach().float()
    return float((a - b).norm() / (b.norm() + 1e-12))


def _build(layers, B, S, seed=1234, lora_b_std=0.02):
    from transformers import MistralConfig, MistralModel
    from oracle.ref_model import LORA_TARGETS, LoraLinear
    from phantom_vlb_b200.engine import Engine
    from phantom_vlb_b200.model import ModelConfig, Trainables
    from phantom_vlb_b200.weights import load_frozen
    torch.manual_seed(seed)
    cfg = ModelConfig(layers=layers, vocab=64, num_target=8, max_len=S)
    mc = MistralConfig(vocab_size=cfg.vocab, hidden_size=cfg.hidden, intermediate_size=cfg.mlp, num_hidden_layers=layers,
                       num_attention_heads=cfg.heads, num_key_value_heads=cfg.kv_heads, head_dim=cfg.head_dim,
                       rms_norm_eps=cfg.rms_eps, rope_theta=cfg.rope_theta, sliding_window=None,
                       max_position_embeddings=4096, use_cache=False)
    mc._attn_implementation = "eager"
    llm = MistralModel(mc)
    for p in llm.parameters():
        p.requires_grad_(False)
    g = torch.Generator().manual_seed(seed + 1)
    for layer in llm.layers:
        for name in LORA_TARGETS:
            parent = layer.self_attn if name.endswith(("q_proj", "k_proj", "v_proj", "o_proj")) else layer.mlp
            lin = LoraLinear(getattr(parent, name), cfg.lora_r, cfg.lora_alpha)
            lin.lora_B.weight.data.copy_(torch.randn(lin.lora_B.weight.shape, generator=g) * lora_b_std)
            setattr(parent, name, lin)
        layer.input_layernorm.weight.data.add_(0.1 * torch.randn(cfg.hidden, generator=g))
        layer.post_attention_layernorm.weight.data.add_(0.1 * torch.randn(cfg.hidden, generator=g))
    with torch.no_grad():
        for p in llm.parameters():
            p.copy_(p.bfloat16().float())
    for n, p in llm.named_parameters():
        p.requires_grad_("lora_" in n)
    dev = torch.device("cuda")
    sd = {"model." + k: v for k, v in llm.state_dict().items() if "lora_" not in k}
    W = load_frozen(sd, cfg, dev, parts=("llm",))
    T = Trainables(cfg, dev)
    H = cfg.hidden
    shim = types.SimpleNamespace(llm=llm, ridge=torch.nn.Linear(H, cfg.num_target), layer_norm1=torch.nn.LayerNorm(H),
                                 layer_norm2=torch.nn.LayerNorm(H))
    T.load_from_oracle(shim)
    return cfg, llm.to(dev), Engine(cfg, W, T)


def test_two_full_width_decoder_layers_fwd_bwd_at_bench_shape():
    B, S = 8, 2048
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    cfg, llm, eng = _build(2, B, S)
    H, M = cfg.hidden, B * S
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(M, H, device="cuda", generator=g).bfloat16()
    gout = (torch.randn(M, H, device="cuda", generator=g) / (M ** 0.5)).bfloat16()     # d loss / d (pre-final-norm hidden)
    # ---- engine: the launches of Engine.loss_and_grads between the splice and the head
    x0 = eng.buf("L0_x", (M, H))
    x0.copy_(x)
    hidden = eng.decoder_forward(x0, B)
    eng._zero_grads()
    dx = eng.buf("bw_dx_a", (M, H))
    dx.copy_(gout)
    eng.decoder_backward(dx, B)
    torch.cuda.synchronize()
    # ---- reference: HF Mistral layers (eager attention, fp32) + restated peft LoRA, same weights, same inputs
    xr = x.float().view(B, S, H)
    with torch.no_grad():
        out = llm(inputs_embeds=xr, attention_mask=torch.ones(B, S, dtype=torch.long, device="cuda"), use_cache=False)
    hp = hidden.float()
    hn = hp * torch.rsqrt(hp.pow(2).mean(-1, keepdim=True) + cfg.rms_eps) * llm.norm.weight
    e_hidden = rel(hn.view(B, S, H), out.last_hidden_state)
    # loss = <pre-final-norm hidden, gout> (what the brain head hands back): run the stack once more without the final norm
    norm = llm.norm
    llm.norm = torch.nn.Identity()
    pre = llm(inputs_embeds=xr, attention_mask=torch.ones(B, S, dtype=torch.long, device="cuda"), use_cache=False).last_hidden_state
    llm.norm = norm
    e_pre = rel(hidden.view(B, S, H), pre)
    (pre * gout.float().view(B, S, H)).sum().backward()
    # context: the error of the reference's OWN arithmetic (bf16 weights and activations through torch) against the same fp32 run
    import copy
    lb = copy.deepcopy(llm).bfloat16()
    lb.norm = torch.nn.Identity()
    with torch.no_grad():
        pre_b = lb(inputs_embeds=x.view(B, S, H), attention_mask=torch.ones(B, S, dtype=torch.long, device="cuda"),
                   use_cache=False).last_hidden_state
    e_torch_bf16 = rel(pre_b, pre)
    del lb, pre_b
    T = eng.T
    gerr = {}
    for l, layer in enumerate(llm.layers):
        a, m = layer.self_attn, layer.mlp
        p = "layers.%d." % l
        ref = {p + "Acat_qkv": torch.cat([a.q_proj.lora_A.weight.grad, a.k_proj.lora_A.weight.grad, a.v_proj.lora_A.weight.grad], 0),
               p + "Bt_q": a.q_proj.lora_B.weight.grad.t(), p + "Bt_k": a.k_proj.lora_B.weight.grad.t(),
               p + "Bt_v": a.v_proj.lora_B.weight.grad.t(), p + "A_o": a.o_proj.lora_A.weight.grad,
               p + "Bt_o": a.o_proj.lora_B.weight.grad.t(),
               p + "Acat_gu": torch.cat([m.gate_proj.lora_A.weight.grad, m.up_proj.lora_A.weight.grad], 0),
               p + "Bt_g": m.gate_proj.lora_B.weight.grad.t(), p + "Bt_u": m.up_proj.lora_B.weight.grad.t(),
               p + "A_d": m.down_proj.lora_A.weight.grad, p + "Bt_d": m.down_proj.lora_B.weight.grad.t()}
        for k, v in ref.items():
            gerr[k] = rel(T.g(k), v)
    print("decoder layers at M=%d: hidden (post-norm) %.3g, pre-norm %.3g; torch bf16 eager on the same weights: %.3g"
          % (M, e_hidden, e_pre, e_torch_bf16))
    print("grad errors:", {k: "%.3g" % v for k, v in gerr.items()})
    assert e_hidden < 1e-2 and e_pre < 1e-2
    worst = max(gerr, key=gerr.get)
    assert gerr[worst] < 3e-2, "gradient %s rel err %g" % (worst, gerr[worst])
