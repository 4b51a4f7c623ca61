"""GPU token parity (-m gpu): BASELINE.json config 4 at the level this repository can pin it.

The reference's end-to-end check is "greedy decode of a random-init MiniCPM4-8B-shaped model produces
identical tokens".  Its model graph (attention, KV cache, EAGLE tree) is out of scope here (SURVEY.md
8 marks it so) and its pybind extension cannot be built on the GPU box, so this test swaps ONLY the
quantised linear — the one thing this library replaces — inside an otherwise identical decoder stack:

    x -> RMSNorm -> qkv (4096->4608) -> [v passthrough: the first 4096 columns stand in for attention]
      -> o (4096->4096) -> +residual -> RMSNorm -> gate_up (4096->32768) -> SiLU(gate)*up
      -> down (16384->4096) -> +residual            (x LAYERS, MiniCPM4-8B layer shapes, group 128)
    logits = RMSNorm(x) @ lm_head^T ; next token = argmax ; next x = embedding[token] + pos[step]

Arm A runs the four linears through the UNMODIFIED reference Marlin kernel (oracle/_ref), arm B
through libb200_w4a16.so (W4A16GPTQMarlinLinear.prefill); every other op is the same torch code.
Greedy tokens must be identical; as SURVEY.md 7 (risk 5) prescribes the requirement is gated on the
reference's own top-1 margin (a near-tie can flip between ANY two correct kernels) and both are logged.
Widths: M = 1 (plain decode) and M = 12 (a tree-verify batch of 12 drafted tokens per step)."""
import pytest
import torch

from oracle import ref_marlin

pytestmark = pytest.mark.gpu
DEV = "cuda"
HID, INTER, QKV, VOCAB = 4096, 16384, 4608, 8192
LAYERS, STEPS = 4, 24
SHAPES = [("qkv", HID, QKV), ("o", HID, HID), ("gate_up", HID, 2 * INTER), ("down", INTER, HID)]


def _rms(x, w):
    # centred: uniformly random uint4b8 weights have mean -0.5, which puts a large common-mode component
    # on every linear's output; without centring the greedy walk collapses onto one token
    xf = x.float()
    xf = xf - xf.mean(-1, keepdim=True)
    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5) * w.float()).half()


def _build(seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    layers = []
    for _ in range(LAYERS):
        lw = {}
        for name, k, n in SHAPES:
            lw[name] = (torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g),
                        (torch.rand((k // 128, n), device=DEV, generator=g) * 0.012 + 0.002).half())
        lw["ln1"] = (1 + 0.1 * torch.randn(HID, device=DEV, generator=g)).half()
        lw["ln2"] = (1 + 0.1 * torch.randn(HID, device=DEV, generator=g)).half()
        layers.append(lw)
    emb = torch.randn((VOCAB, HID), device=DEV, generator=g).half()
    head = (torch.randn((VOCAB, HID), device=DEV, generator=g) * 0.05).half()
    lnf = torch.ones(HID, device=DEV).half()
    pos = (0.7 * torch.randn((STEPS, HID), device=DEV, generator=g)).half()  # keeps the greedy walk from a fixed point
    return layers, emb, head, lnf, pos


class _RefArm:
    def __init__(self, m):
        self.ws = torch.zeros((2 * INTER // 64) * 16, dtype=torch.int32, device=DEV)
        self.ct = torch.empty(ref_marlin.reduce_max_m(m) * 2 * INTER, dtype=torch.float32, device=DEV)

    def linear(self, x, w):
        return ref_marlin.gemm(x.contiguous(), w[0], w[1], self.ws, self.ct)


class _OurArm:
    def __init__(self, m):
        import cpmcu_b200 as B
        self.B = B
        self.ws = torch.zeros(B.workspace_ints(2 * INTER), dtype=torch.int32, device=DEV)
        self.ct = torch.empty(max(B.reduce_bytes_fast(m, n, k) for _, k, n in SHAPES) // 4 + 64, dtype=torch.float32,
                              device=DEV)

    def linear(self, x, w):
        return self.B.gemm(x.contiguous(), w[0], w[1], self.ws, self.ct, pdl=True)


def _decode(arm, layers, emb, head, lnf, pos, first_tokens):
    tok = first_tokens.clone()
    out_tokens, margins = [], []
    for step in range(STEPS):
        x = (emb[tok].float() + pos[step].float()).half()
        for lw in layers:
            h = _rms(x, lw["ln1"])
            qkv = arm.linear(h, lw["qkv"])
            att = qkv[:, :HID]                      # stand-in for attention: same torch op in both arms
            x = (x.float() + 0.05 * arm.linear(att, lw["o"]).float()).half()   # scale_residual-style damping
            h = _rms(x, lw["ln2"])
            gu = arm.linear(h, lw["gate_up"])
            act = (torch.nn.functional.silu(gu[:, :INTER].float()) * gu[:, INTER:].float()).half()
            x = (x.float() + 0.01 * arm.linear(act, lw["down"]).float()).half()
        logits = _rms(x, lnf).float() @ head.float().t()
        top2 = logits.topk(2, dim=-1).values
        margins.append((top2[:, 0] - top2[:, 1]).cpu())
        tok = logits.argmax(-1)
        out_tokens.append(tok.cpu())
    return torch.stack(out_tokens), torch.stack(margins)


@pytest.mark.parametrize("m", [1, 12])
def test_greedy_tokens_match_reference_marlin(m):
    if not ref_marlin.available():
        pytest.skip("oracle/_ref/libmarlin_ref.so not built")
    layers, emb, head, lnf, pos = _build(seed=100 + m)
    first = torch.arange(m, device=DEV) * 37 % VOCAB
    ref_tok, ref_margin = _decode(_RefArm(m), layers, emb, head, lnf, pos, first)
    our_tok, _ = _decode(_OurArm(m), layers, emb, head, lnf, pos, first)
    torch.cuda.synchronize()
    same = (ref_tok == our_tok)
    # a mismatch is only admissible where the reference's own top-1 margin is below what one fp16 ulp of
    # the hidden state can move a logit (~1e-2 here); after a legitimately flipped token the sequences
    # diverge, so only the prefix up to the first mismatch of each stream is compared
    MARGIN = 2e-2
    mism = 0
    for j in range(m):
        bad = (~same[:, j]).nonzero().flatten()
        if len(bad):
            t = int(bad[0])
            mism += 1
            assert float(ref_margin[t, j]) < MARGIN, (
                f"stream {j}: token {t} differs ({int(ref_tok[t, j])} vs {int(our_tok[t, j])}) "
                f"although the reference margin is {float(ref_margin[t, j]):.4f}")
    print(f"\n[token parity m={m}] {STEPS} greedy steps x {m} streams, {LAYERS} layers: "
          f"{int(same.all(0).sum())}/{m} streams identical end to end, {mism} near-tie flips; "
          f"reference top-1 margin min {float(ref_margin.min()):.4f} median {float(ref_margin.median()):.4f}; "
          f"{len(set(ref_tok.flatten().tolist()))} distinct tokens")
    assert mism <= max(1, m // 6), "too many near-tie flips: the kernels drift"
    assert len(set(ref_tok.flatten().tolist())) >= STEPS // 2, "degenerate greedy walk: the test would prove nothing"
