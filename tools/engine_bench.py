"""End-to-end numbers through `LLM` at BASELINE.json's model shapes with synthetic weights:

    python tools/engine_bench.py --model llama-3.1-8b --prompt 16384 --ratio 0.5 --method COMPACTOR --new 256
    python tools/engine_bench.py --model qwen3-8b --prompt 40960 --ratio 0.25 --method SNAPKV --batch 16 --new 64

Prints one JSON line: time to first token (prefill incl. scoring / compaction / reclaim), prefill tok/s,
decode tok/s from CUDA-graph replays, pages in use after the prefill.  Not bench.py's headline (that is the
attention + compression path alone); this is the same path inside a whole model step, where the cuBLAS
projections and the MLP take most of the time."""
import argparse
import json
import sys
import time
from pathlib import Path
from types import SimpleNamespace

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]

SHAPES = {
    "llama-3.1-8b": dict(model_type="llama", vocab_size=128256, hidden_size=4096, intermediate_size=14336,
                         num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=8, head_dim=128,
                         rms_norm_eps=1e-5, rope_theta=500000.0, max_position_embeddings=131072,
                         rope_scaling={"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0,
                                       "high_freq_factor": 4.0, "original_max_position_embeddings": 8192}),
    "qwen3-8b": dict(model_type="qwen3", vocab_size=151936, hidden_size=4096, intermediate_size=12288,
                     num_hidden_layers=36, num_attention_heads=32, num_key_value_heads=8, head_dim=128,
                     rms_norm_eps=1e-6, rope_theta=1000000.0, max_position_embeddings=40960, rope_scaling=None),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-3.1-8b", choices=list(SHAPES))
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (0 = the model's)")
    ap.add_argument("--prompt", type=int, default=16384)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--ratio", type=float, default=0.5)
    ap.add_argument("--method", default="COMPACTOR", choices=["COMPACTOR", "SNAPKV", "NONE"])
    ap.add_argument("--new", type=int, default=256)
    ap.add_argument("--eager", action="store_true")
    a = ap.parse_args()

    from compactor_vllm import LLM, CompressionMethod, LLMConfig, SamplingParams
    from compactor_vllm.api_types import BatchCompressionParams, SequenceCompressionParams
    from compactor_vllm.utils.sequence import Sequence, SequenceStatus

    shape = dict(SHAPES[a.model])
    if a.layers:
        shape["num_hidden_layers"] = a.layers
    hf = SimpleNamespace(torch_dtype=torch.bfloat16, tie_word_embeddings=False, attention_bias=False, **shape)
    max_len = a.prompt + a.new + 128
    cfg = LLMConfig(model="synthetic", hf_config=hf, max_num_seqs=max(a.batch, 1), max_model_len=max_len,
                    kvcache_page_size=128, eos=-1, enforce_eager=a.eager)
    pages_per_seq = hf.num_key_value_heads * ((max_len + 127) // 128 + 1)
    t0 = time.perf_counter()
    llm = LLM(cfg, num_pages=pages_per_seq * (a.batch + 1), max_batched_tokens=max(a.prompt, 32768))
    torch.cuda.synchronize()
    t_init = time.perf_counter() - t0
    run = llm.model_runner
    g = torch.Generator().manual_seed(0)
    bp = BatchCompressionParams(compression_method=CompressionMethod[a.method])
    cp = SequenceCompressionParams(compression_ratio=a.ratio if a.method != "NONE" else 1.0)
    sp = SamplingParams(temperature=0.0, max_new_tokens=a.new)

    def fresh():
        return [Sequence(prompt_token_ids=torch.randint(0, hf.vocab_size, (a.prompt,), generator=g).tolist(),
                         sampling_params=sp,
                         compression_params=SequenceCompressionParams(cp.compression_ratio, cp.protected_first_tokens,
                                                                      cp.protected_last_tokens))
                for _ in range(a.batch)]

    def one_pass(seqs):
        torch.cuda.synchronize()
        t = time.perf_counter()
        for batch in run._prefill_batches(seqs):
            logits = run.run_prefill(batch, bp)
            first = logits.float().argmax(-1).tolist()
            for s, tok in zip(batch, first):
                s.status = SequenceStatus.RUNNING
                run._accept(s, tok)
        torch.cuda.synchronize()
        t_prefill = time.perf_counter() - t
        used = int(run.kv.paged_cache.bh_num_pages.sum())
        t = time.perf_counter()
        run.run_decode_loop(seqs)
        torch.cuda.synchronize()
        t_decode = time.perf_counter() - t
        run.kv.free_sequences([s.seq_id for s in seqs])
        return t_prefill, t_decode, used

    one_pass(fresh())  # warm-up: workspaces, graph capture, cuBLAS heuristics
    steps0 = run.stats["decode_steps"]
    t_prefill, t_decode, used = one_pass(fresh())
    steps = run.stats["decode_steps"] - steps0
    full_pages = a.batch * hf.num_hidden_layers * hf.num_key_value_heads * ((a.prompt + 127) // 128)
    print(json.dumps({
        "model": a.model, "layers": hf.num_hidden_layers, "weights": "synthetic", "prompt_tokens": a.prompt,
        "batch": a.batch, "method": a.method, "ratio": cp.compression_ratio, "new_tokens": a.new,
        "ttft_s": t_prefill, "prefill_tok_s": a.batch * a.prompt / t_prefill,
        "decode_tok_s": a.batch * steps / t_decode, "decode_ms_per_step": 1e3 * t_decode / max(steps, 1),
        "decode": "eager" if a.eager else "cuda graph", "kv_pages_after_prefill": used,
        "kv_pages_uncompressed": full_pages, "init_s": t_init,
    }))


if __name__ == "__main__":
    main()
