#!/usr/bin/env python
"""TEST INFRASTRUCTURE (BASELINE.json configs[3]): drives the REFERENCE's own model graph — its pybind extension
`cpmcu.C` built by `make -C oracle cpmcu` (src/entry.cu:193-235,359-407,536-570) — with a random-init
MiniCPM4-8B-shaped W4A16 model, greedy decoding, optionally with the EAGLE / FR-Spec draft tree.

Two builds of the SAME objects exist under oracle/_ref/:
    cpmcu_vanilla/C.so   linked with the reference's Marlin kernels
    cpmcu_b200/C.so      linked with include/gptq_marlin_gemm_shim.cuh + libb200_w4a16.so (nothing else differs)
This script runs ONE of them in its own process (the extension keeps a global model and owns most of the GPU's
memory) and writes tokens, per-step top-2 logits and timings to a JSON file; tests/test_gpu_cpmcu_tokens.py compares
the two.  The call sequence mirrors cpmcu/llm_w4a16_gptq_marlin.py:71-264 (init, init_storage, load_model per tensor,
chunked prefill, decode with the CUDA-graph flag) and cpmcu/speculative/tree_drafter_base_quant/
tree_drafter_w4a16_gptq_marlin.py:183-214 (draft -> tree decode -> verify_and_fix).  Nothing here reads /root/reference.

    python tests/cpmcu_driver.py --build vanilla|b200 --out x.json [--eagle] [--layers 32] [--prompt 1500] [--gen 64]
                                 [--force-tokens y.json]   # teacher forcing: feed y's tokens, record own argmax
"""
import argparse
import importlib.util
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# MiniCPM4-8B (public config): scale_emb 12, dim_model_base 256, scale_depth 1.4; EAGLE layer of the same width
CFG = dict(vocab=73448, hidden=4096, inter=16384, heads=32, kv_heads=2, head_dim=128, eps=1e-5, group=128,
           scale_emb=12.0, dim_model_base=256, scale_depth=1.4, rope_theta=10000.0)
EAGLE = dict(layers=1, num_iter=2, topk=10, tree=12, window=1024, frspec=32768)


def load_ext(build):
    path = os.path.join(ROOT, "oracle", "_ref", f"cpmcu_{build}", "C.so")
    spec = importlib.util.spec_from_file_location("C", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--build", required=True, choices=["vanilla", "b200"])
    ap.add_argument("--out", required=True)
    ap.add_argument("--eagle", action="store_true")
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--prompt", type=int, default=1500)
    ap.add_argument("--gen", type=int, default=64)
    ap.add_argument("--chunk", type=int, default=1024)
    ap.add_argument("--memory-limit", type=float, default=0.25)
    ap.add_argument("--force-tokens", default="")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--fast-scratch", type=int, default=1,
                    help="b200 build: declare the model's real c_tmp size (b200_w4a16_set_scratch_capacity), as an integration would")
    args = ap.parse_args()

    import torch
    C = load_ext(args.build)
    dev = "cuda"
    c = CFG
    L = args.layers
    scale_residual = c["scale_depth"] / math.sqrt(L)
    if args.build == "b200":
        import cpmcu_b200 as B
        B.lib()
        B.set_pdl(True)
        if args.fast_scratch:
            # what the model allocates: reduce_max_m(chunk) * 2*intermediate floats per layer scratch
            # (w4a16_gptq_marlin_layer.cuh:34-37) and reduce_max_m(chunk) * 2*hidden floats for EAGLE (minicpm4_eagle.cuh:134-137)
            B.set_scratch_capacity(B.reduce_max_m(args.chunk) * 2 * c["hidden"] * 4)

    C.init_w4a16_gptq_marlin_base_model(args.memory_limit, c["vocab"], L, c["hidden"], c["inter"], c["heads"], c["kv_heads"],
                                        c["head_dim"], c["eps"], c["group"], 0, args.chunk, c["scale_emb"],
                                        c["dim_model_base"] / c["hidden"], scale_residual, False, False)
    if args.eagle:
        C.init_minicpm4_eagle_model(EAGLE["layers"], c["inter"], c["heads"], c["kv_heads"], c["head_dim"], c["eps"],
                                    EAGLE["num_iter"], EAGLE["topk"], EAGLE["tree"], 0, True, c["group"], EAGLE["window"],
                                    EAGLE["frspec"], c["scale_depth"] / math.sqrt(L + 1), True, True)
    max_total = C.init_storage()

    # ---- random-init weights, identical in both builds (seeded per tensor); C.load_model takes HOST pointers ----
    def seeded(name):
        g = torch.Generator(device=dev)
        g.manual_seed(abs(hash_name(name)) % (2**31))
        return g

    def hash_name(name):
        h = 1469598103934665603
        for ch in name.encode():
            h = ((h ^ ch) * 1099511628211) & ((1 << 63) - 1)
        return h

    def load(name, t):
        t = t.contiguous().cpu()
        C.load_model(name, t.data_ptr())

    def qlinear(prefix, k, n, bias=False):
        g = seeded(prefix)
        q = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
        # no zero nibbles: a nibble 0 becomes 8, so that E[q - 8] = 0 (no common-mode component in every output)
        z = ~(q | (q >> 1) | (q >> 2) | (q >> 3)) & 0x11111111
        q = q | (z << 3)
        s = (torch.rand((k // 128, n), device=dev, generator=g) * 0.006 + 0.002).half()
        load(prefix + ".qweight", q)
        load(prefix + ".scales", s)
        if bias:
            load(prefix + ".bias", (torch.randn(n, device=dev, generator=g) * 0.02).half())

    def norm(name):
        load(name, (1.0 + 0.1 * torch.randn(c["hidden"], device=dev, generator=seeded(name))).half())

    qkv_n = (c["heads"] + 2 * c["kv_heads"]) * c["head_dim"]
    inv_freq = 1.0 / (c["rope_theta"] ** (torch.arange(0, c["head_dim"], 2, dtype=torch.float32) / c["head_dim"]))

    def layer(prefix):
        qlinear(prefix + ".self_attn.qkv_proj", c["hidden"], qkv_n)
        qlinear(prefix + ".self_attn.o_proj", c["heads"] * c["head_dim"], c["hidden"])
        qlinear(prefix + ".mlp.gate_up_proj", c["hidden"], 2 * c["inter"])
        qlinear(prefix + ".mlp.down_proj", c["inter"], c["hidden"])
        norm(prefix + ".input_layernorm.weight")
        norm(prefix + ".post_attention_layernorm.weight")

    t0 = time.time()
    if args.eagle:
        # drafter tensors first, the FR-Spec remap before lm_head (minicpm4_eagle.cuh:192-226 copies the remapped head at lm_head load)
        remap = torch.randperm(c["vocab"], generator=torch.Generator().manual_seed(5))[:EAGLE["frspec"]].sort().values.to(torch.int32)
        load("eagle.token_id_remap", remap)
        qlinear("eagle.fc1", c["hidden"], c["hidden"], bias=True)
        qlinear("eagle.fc2", c["hidden"], c["hidden"])
        norm("eagle.input_norm1.weight")
        norm("eagle.input_norm2.weight")
        layer("eagle.layers.0")
        load("eagle.rotary_emb.inv_freq", inv_freq)
    load("model.embed_tokens.weight", (torch.randn((c["vocab"], c["hidden"]), device=dev, generator=seeded("embed")) * 0.05).half())
    for i in range(L):
        layer(f"model.layers.{i}")
    norm("model.norm.weight")
    # wide top-1 margins (SURVEY.md section 7 risk 5): a large head scale spreads the logits
    load("lm_head.weight", (torch.randn((c["vocab"], c["hidden"]), device=dev, generator=seeded("lm_head")) * 2.0).half())
    load("model.rotary_emb.inv_freq", inv_freq)
    torch.cuda.synchronize()
    t_load = time.time() - t0

    logits = torch.empty((64, c["vocab"]), dtype=torch.float16, device=dev)
    prompt = torch.randint(0, c["vocab"], (args.prompt,), dtype=torch.int32, device=dev,
                           generator=torch.Generator(device=dev).manual_seed(7))
    pos = torch.arange(args.prompt, dtype=torch.int32, device=dev)
    assert args.prompt + args.gen + 64 <= max_total, (args.prompt, max_total)

    def top2(row):
        v, i = torch.topk(row.float(), 2)
        return [float(v[0]), float(v[1]), int(i[0]), int(i[1])]

    forced = json.load(open(args.force_tokens))["tokens"] if args.force_tokens else None
    graph = not args.no_graph

    def generate():
        # ---- chunked prefill (llm_w4a16_gptq_marlin.py:203-242) ----
        torch.cuda.synchronize()
        t0 = time.time()
        for i in range(0, args.prompt, args.chunk):
            n = min(args.prompt - i, args.chunk)
            C.prefill(n, i, prompt[i:].data_ptr(), pos[i:].data_ptr(), logits.data_ptr())
        torch.cuda.synchronize()
        t_prefill = time.time() - t0
        steps = [dict(kind="prefill", top2=top2(logits[0]))]
        first = int(logits[0].argmax())
        tokens = [forced[0] if forced else first]
        own = [first]

        def decode(ids, position_ids, cache_length, mask=None):  # llm_w4a16_gptq_marlin.py:244-264
            n = ids.numel()
            cache_length += n
            padded = (int(cache_length[0].item()) + 127) // 128 * 128
            C.decode(n, padded, ids.data_ptr(), position_ids.data_ptr(), cache_length.data_ptr(),
                     mask.data_ptr() if mask is not None else 0, logits.data_ptr(), graph)
            cache_length -= n
            return logits[:n]

        accept_lengths = []
        torch.cuda.synchronize()
        t0 = time.time()
        if not args.eagle:
            ids = torch.zeros(1, dtype=torch.int32, device=dev)
            p1 = torch.zeros(1, dtype=torch.int32, device=dev)
            cl = torch.zeros(1, dtype=torch.int32, device=dev)
            for i in range(args.gen - 1):
                ids[0] = tokens[-1]
                p1[0] = args.prompt + i
                cl[0] = args.prompt + i
                lg = decode(ids, p1, cl)
                steps.append(dict(kind="decode", top2=top2(lg[0])))
                nxt = int(lg[0].argmax())
                own.append(nxt)
                tokens.append(forced[i + 1] if forced else nxt)
        else:
            T = EAGLE["tree"]
            tree_ids = torch.zeros(T, dtype=torch.int32, device=dev)
            tree_pos = torch.zeros(T, dtype=torch.int32, device=dev)
            tree_gt = torch.zeros(T, dtype=torch.int32, device=dev)
            tree_mask = torch.zeros(T, dtype=torch.int64, device=dev)
            tree_parent = torch.zeros(T, dtype=torch.int32, device=dev)
            cl = torch.zeros(1, dtype=torch.int32, device=dev)
            tree_ids[0] = tokens[0]
            i = 0
            while i < args.gen - 1:
                cl[0] = args.prompt + i
                C.draft(tree_ids.data_ptr(), tree_pos.data_ptr(), cl.data_ptr(), tree_mask.data_ptr(), tree_parent.data_ptr())
                lg = decode(tree_ids, tree_pos, cl, mask=tree_mask)
                tree_gt.copy_(lg.argmax(dim=-1))
                rows = [top2(lg[r]) for r in range(T)]
                acc = C.verify_and_fix(T, tree_ids.data_ptr(), tree_gt.data_ptr(), tree_pos.data_ptr(), cl.data_ptr(),
                                       tree_mask.data_ptr(), tree_parent.data_ptr())
                accept_lengths.append(int(acc))
                app = min(acc, args.gen - 1 - i)
                new = tree_ids[:app].tolist()
                tokens.extend(int(t) for t in new)
                own.extend(int(t) for t in new)
                steps.append(dict(kind="tree", accept=int(acc), min_margin=min(r[0] - r[1] for r in rows), top2=rows[0]))
                tree_ids[0] = tree_ids[acc - 1]
                i += acc
        torch.cuda.synchronize()
        t_decode = time.time() - t0
        return tokens, own, steps, accept_lengths, t_prefill, t_decode

    generate()   # warm-up pass: module load, descriptor encoding, CUDA-graph capture; same prompt, same KV positions
    tokens, own, steps, accept_lengths, t_prefill, t_decode = generate()

    ngen = len(tokens) - 1
    out = dict(build=args.build, eagle=bool(args.eagle), layers=L, prompt=args.prompt, gen=args.gen, tokens=tokens, own_argmax=own,
               steps=steps, accept_lengths=accept_lengths, max_total_length=int(max_total),
               seconds=dict(load=round(t_load, 2), prefill=round(t_prefill, 4), decode=round(t_decode, 4)),
               prefill_tok_s=round(args.prompt / t_prefill, 1), decode_tok_s=round(ngen / t_decode, 2),
               cuda_graph=graph, forced=bool(forced))
    if args.build == "b200":
        out["b200_launches"] = int(B.launch_count())
    json.dump(out, open(args.out, "w"))
    print(f"[{args.build}{'+eagle' if args.eagle else ''}] {ngen} tokens, prefill {out['prefill_tok_s']} tok/s, decode "
          f"{out['decode_tok_s']} tok/s, mean accept {sum(accept_lengths) / max(1, len(accept_lengths)):.2f}", flush=True)
    os._exit(0)   # the extension installs signal handlers and owns a huge pool: leave without teardown


if __name__ == "__main__":
    main()
