"""Informational (GPU): decode throughput of vLLM (the engine family the reference deploys, installed
in the image) on the bench workload C2 -- batch 64, 512-token prompts, 256 new tokens, Qwen3-0.6B
architecture with dummy (random) weights.  Not the graded denominator; it
says how the stock stack does on the same box.  Decode tok/s = 64 * 255 / (t(256 tokens) - t(1 token))."""
import json, os, sys, tempfile, time

def main():
    d = tempfile.mkdtemp(prefix="qwen3_cfg_")
    cfg = {"architectures": ["Qwen3ForCausalLM"], "model_type": "qwen3", "hidden_size": 1024, "intermediate_size": 3072,
           "num_hidden_layers": 28, "num_attention_heads": 16, "num_key_value_heads": 8, "head_dim": 128, "vocab_size": 151936,
           "max_position_embeddings": 40960, "rms_norm_eps": 1e-6, "rope_theta": 1000000.0, "tie_word_embeddings": True,
           "hidden_act": "silu", "attention_bias": False, "torch_dtype": "bfloat16", "bos_token_id": 151643, "eos_token_id": 151645}
    with open(os.path.join(d, "config.json"), "w") as f:
        json.dump(cfg, f)
    from vllm import LLM, SamplingParams
    import numpy as np
    llm = LLM(model=d, load_format="dummy", skip_tokenizer_init=True, dtype="bfloat16", max_model_len=1024,
              gpu_memory_utilization=0.5, max_num_seqs=64, max_num_batched_tokens=8192, seed=0, enable_prefix_caching=False)
    rng = np.random.default_rng(1)
    prompts = [{"prompt_token_ids": rng.integers(0, 151936, 512).tolist()} for _ in range(64)]
    def run(n):
        sp = SamplingParams(temperature=0.0, max_tokens=n, ignore_eos=True, detokenize=False)
        t0 = time.perf_counter()
        out = llm.generate(prompts, sp, use_tqdm=False)
        dt = time.perf_counter() - t0
        assert all(len(o.outputs[0].token_ids) == n for o in out)
        return dt
    run(8)  # warm-up (graph capture, allocator)
    t1 = min(run(1) for _ in range(2))
    t256 = min(run(256) for _ in range(2))
    import vllm
    print("VLLM_RESULT " + json.dumps({"vllm": vllm.__version__, "t_1tok_s": t1, "t_256tok_s": t256,
                                       "decode_tok_s": 64 * 255 / (t256 - t1), "ms_per_step": (t256 - t1) / 255 * 1e3}), flush=True)

if __name__ == "__main__":
    main()
