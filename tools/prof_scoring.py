"""One pass over every scoring / selection / store kernel at the headline shape (32768 tokens, HQ=32 HKV=8 D=128),
for ncu: COMPACTOR chain (leverage_*, chunk_scores, sel_*), SnapKV (snap_stats / snap_scores), store-all.

    python tools/prof_scoring.py            # 3 warm-up passes + 1 pass; prints per-stage CUDA-event times
"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    from compactor_vllm.compression.snapkv import query_aware_key_scores
    from compactor_vllm.kv_cache.store_kv_cache import prefill_store_all_kv
    w = bench.Workload(dev, bench.SEQ, seed=1234)

    def snap():
        return query_aware_key_scores(w.q, w.k, w.cu, w.cu, w=32)

    def store_all():
        w.reset_slot()
        lens = w.cache.bh_seq_lens[0].index_select(0, w.batch_mapping).contiguous()
        prefill_store_all_kv(new_keys=w.k, new_values=w.v, cu_seqlens_k=w.cu, max_seqlen_k=w.n_tok,
                             k_cache=w.attn.k_cache, v_cache=w.attn.v_cache, page_table=w.attn.page_table,
                             bh_lens=lens, batch_mapping=w.batch_mapping, PAGE_SIZE=bench.PS)

    from compactor_vllm.compression.compactor import non_causal_attn_scores

    def chunk_fat():
        return non_causal_attn_scores(w.q, w.k, w.v, w.cu, w.n_tok, 128, sm_scale=1.0, normalize=True)

    stages = {"compactor_chain": w.compress_only, "chunk_scores(+finalize)": chunk_fat, "snapkv": snap,
              "store_all": store_all}
    for name, fn in stages.items():
        try:
            once = "--once" in sys.argv  # under ncu --set full: one warm-up pass + one pass per stage
            for _ in range(1 if once else 3):
                fn()
            torch.cuda.synchronize()
            print(f"{name}: {bench.time_loop(fn, 1 if once else 5) * 1e3:.1f} us", flush=True)
        except Exception as ex:  # a signature drift must not hide the other stages from ncu
            print(f"{name}: FAILED {ex!r}", flush=True)
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
