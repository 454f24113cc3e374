"""Engine shell: `LLM.generate` drives the hot path end to end — prefill with scoring and
compaction on the store stream, page reclaim, CUDA-graph decode — on a small synthetic-weight model.

Parity: without compression the engine's greedy tokens are checked against a plain fp32 torch
restatement of the same decoder (oracle RMSNorm / RoPE, dense causal softmax attention) by teacher
forcing: at every step the engine's token must be the reference's argmax up to a logit margin that
covers bf16 rounding.  Graph replay must equal eager decode token for token."""
import math
from types import SimpleNamespace

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0) if torch.cuda.is_available() else None


def tiny_config(model_type="llama", layers=2, **over):
    from compactor_vllm import LLMConfig

    hf = SimpleNamespace(
        model_type=model_type, vocab_size=384, hidden_size=512, intermediate_size=1024, num_hidden_layers=layers,
        num_attention_heads=4, num_key_value_heads=2, head_dim=128, rms_norm_eps=1e-5, rope_theta=500000.0,
        rope_scaling=None, max_position_embeddings=4096, torch_dtype=torch.bfloat16, tie_word_embeddings=False,
        attention_bias=False)
    kw = dict(model="synthetic", hf_config=hf, max_num_seqs=8, max_model_len=4096, kvcache_page_size=128, eos=-1)
    kw.update(over)
    return LLMConfig(**kw)


def make_llm(cfg, **kw):
    from compactor_vllm import LLM

    llm = LLM(cfg, num_pages=kw.pop("num_pages", 512), **kw)
    with torch.no_grad():  # sharper logits: the argmax of a random model is otherwise a near-tie
        llm.model_runner.model.lm_head.weight.mul_(6.0)
    return llm


def prompts(lens, vocab=384, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randint(0, vocab, (n,), generator=g).tolist() for n in lens]


@torch.no_grad()
def reference_logits(model, token_ids):
    """fp32 CPU restatement of models/decoder.py for ONE sequence, dense causal attention: logits [T, vocab]."""
    from oracle import prepath as op

    sd = {k: v.detach().float().cpu() for k, v in model.state_dict().items()}
    cfg = model.config
    H, HKV, D = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
    T = len(token_ids)
    cache = op.cos_sin_cache(D, T, cfg.rope_theta, None)
    pos = torch.arange(T)
    bf = lambda t: t.to(torch.bfloat16)
    h = sd["model.embed_tokens.weight"][torch.tensor(token_ids)]
    residual = None
    for l in range(cfg.num_hidden_layers):
        pre = f"model.layers.{l}."
        if residual is None:
            residual = h
            x = op.rms_norm(bf(h), bf(sd[pre + "input_layernorm.weight"]), cfg.rms_norm_eps).float()
        else:
            x, r = op.rms_norm(bf(h), bf(sd[pre + "input_layernorm.weight"]), cfg.rms_norm_eps, bf(residual))
            x, residual = x.float(), r.float()
        qkv = x @ sd[pre + "self_attn.qkv_proj.weight"].T
        q, k, v = qkv.split([H * D, HKV * D, HKV * D], -1)
        q, k, v = q.view(T, H, D), k.view(T, HKV, D), v.view(T, HKV, D)
        if pre + "self_attn.q_norm.weight" in sd:
            q = op.rms_norm(bf(q), bf(sd[pre + "self_attn.q_norm.weight"]), cfg.rms_norm_eps).float()
            k = op.rms_norm(bf(k), bf(sd[pre + "self_attn.k_norm.weight"]), cfg.rms_norm_eps).float()
        q = op.rope(bf(q), pos, cache).float()
        k = op.rope(bf(k), pos, cache).float()
        kk = k.repeat_interleave(H // HKV, dim=1)
        vv = v.repeat_interleave(H // HKV, dim=1)
        s = torch.einsum("thd,shd->hts", q, kk) / math.sqrt(D)
        s = s.masked_fill(torch.triu(torch.ones(T, T, dtype=torch.bool), 1), float("-inf"))
        o = torch.einsum("hts,shd->thd", torch.softmax(s, -1), vv).reshape(T, H * D)
        a = o @ sd[pre + "self_attn.o_proj.weight"].T
        x, r = op.rms_norm(bf(a), bf(sd[pre + "post_attention_layernorm.weight"]), cfg.rms_norm_eps, bf(residual))
        x, residual = x.float(), r.float()
        gate, up = (x @ sd[pre + "mlp.gate_up_proj.weight"].T).chunk(2, -1)
        h = (torch.nn.functional.silu(gate) * up) @ sd[pre + "mlp.down_proj.weight"].T
    x, _ = op.rms_norm(bf(h), bf(sd["model.norm.weight"]), cfg.rms_norm_eps, bf(residual))
    return x.float() @ sd["lm_head.weight"].T


@pytest.mark.parametrize("model_type", ["llama", "qwen3"])
def test_generate_matches_dense_reference_without_compression(model_type):
    from compactor_vllm import SamplingParams

    llm = make_llm(tiny_config(model_type))
    ps = prompts([37, 300, 129])
    outs, seqs = llm.generate(ps, SamplingParams(temperature=0.0, max_new_tokens=12), return_sequences=True)
    assert [len(o) for o in outs] == [12, 12, 12]
    assert llm.model_runner.stats["graph_replays"] == 11
    for p, out in zip(ps, outs):
        logits = reference_logits(llm.model_runner.model, p + out)
        for i, tok in enumerate(out):
            row = logits[len(p) + i - 1]
            assert row.max().item() - row[tok].item() <= 0.25, (i, tok, int(row.argmax()))
    # all pages came back
    assert llm.model_runner.kv.num_free_pages == 512


def test_graph_decode_equals_eager_decode():
    from compactor_vllm import SamplingParams

    ps = prompts([64, 513, 200, 90, 31], seed=3)
    sp = SamplingParams(temperature=0.0, max_new_tokens=20)
    a = make_llm(tiny_config(enforce_eager=False)).generate(ps, sp)
    b = make_llm(tiny_config(enforce_eager=True)).generate(ps, sp)
    assert a == b


@pytest.mark.parametrize("method,model_type", [("COMPACTOR", "llama"), ("SNAPKV", "qwen3"), ("COMPACTOR", "qwen3")])
def test_generate_with_compression(method, model_type):
    """Compressed prefill: the cache holds the retain budget rounded up to whole pages per head, the freed
    pages go back to the allocator, and decoding on the compacted cache still produces tokens."""
    from compactor_vllm import CompressionMethod, SamplingParams
    from compactor_vllm.api_types import BatchCompressionParams, SequenceCompressionParams
    from compactor_vllm.utils.sequence import Sequence

    llm = make_llm(tiny_config(model_type))
    run = llm.model_runner
    bp = BatchCompressionParams(compression_method=CompressionMethod[method])
    ps = prompts([1500, 700], seed=5)
    cp = SequenceCompressionParams(compression_ratio=0.25, protected_first_tokens=16, protected_last_tokens=64)
    # prefill alone first, to look at the cache
    seqs = [Sequence(prompt_token_ids=p, compression_params=cp) for p in ps]
    free0 = run.kv.num_free_pages
    run.run_prefill(seqs, bp)
    torch.cuda.synchronize()
    HKV, PS = 2, 128
    for s in seqs:
        slot = run.kv.seq_id_to_batch[s.seq_id]
        lens = run.kv.paged_cache.bh_seq_lens[:, slot].cpu()  # [layers, HKV]
        retain = max(round(0.25 * (s.num_prompt_tokens - 80) * HKV), 1)
        for l in range(lens.shape[0]):
            kept = int(lens[l].sum())
            assert retain <= kept <= retain + HKV * (PS - 1), (l, kept, retain)
            assert (lens[l] <= s.num_prompt_tokens).all()
        pages = run.kv.paged_cache.bh_num_pages[:, slot].cpu()
        assert torch.equal(pages, (lens + PS - 1) // PS)  # reclaimed down to what is used
    assert run.stats["bytes_reclaimed"] > 0 and run.kv.num_free_pages > free0 - (12 + 6) * HKV
    run.kv.free_sequences([s.seq_id for s in seqs])
    assert run.kv.num_free_pages == free0
    # whole call through the public API
    outs = llm.generate(ps, SamplingParams(temperature=0.0, max_new_tokens=8), bp, per_sequence_compression_params=cp)
    assert [len(o) for o in outs] == [8, 8] and all(0 <= t < 384 for o in outs for t in o)
    assert run.kv.num_free_pages == free0


def test_sampling_temperature_and_eos():
    from compactor_vllm import SamplingParams

    llm = make_llm(tiny_config())
    ps = prompts([50, 60], seed=9)
    greedy = llm.generate(ps, SamplingParams(temperature=0.0, max_new_tokens=6))
    torch.manual_seed(0)
    hot = llm.generate(ps, [SamplingParams(temperature=5.0, max_new_tokens=6), SamplingParams(temperature=0.0, max_new_tokens=6)])
    assert hot[1] == greedy[1]            # temperature 0 rows stay greedy inside a mixed batch
    assert hot[0] != greedy[0]            # 6 draws at temperature 5 over 384 tokens
    # eos stops a sequence early; the other keeps decoding with the finished row parked on the reserved slot
    llm.model_runner.eos = greedy[0][2]
    out = llm.generate(ps, SamplingParams(temperature=0.0, max_new_tokens=6))
    stop = greedy[0].index(greedy[0][2]) + 1  # the eos token itself is kept, as in the reference
    assert out[0] == greedy[0][:stop] and out[1] == greedy[1][:len(out[1])] and len(out[1]) >= 1
