#!/usr/bin/env python
"""bf16 error against the fp32 oracle as a function of sequence length (P is rounded to bf16 before
P.V -- watch the 1e-2 budget at long S).  H=8, D=128, B=1, causal and non-causal, standard-normal q/k/v; the oracle is
evaluated on a sample of rows (all rows up to S=2048).  Reports max-abs and mean-abs error of ours, and -- where the
libraries are importable -- of the reference's own kernels (oracle/_ref) and of flash-attn 2.8.3 on the same inputs.
One JSON line per case; run on the GPU box, keep the output under profiles/."""
import importlib.util, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import mini_flash_attention as mfa


def load_ref():
    so = os.path.join(ROOT, "oracle", "_ref", "mini_flash_attention_ref_C.so")
    if not os.path.exists(so):
        return None
    spec = importlib.util.spec_from_file_location("mini_flash_attention_ref_C", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def rows_oracle(q, k, v, rows, causal):
    D, S = q.shape[-1], k.shape[1]
    qs = q[:, rows].float().permute(0, 2, 1, 3)
    s = torch.matmul(qs, k.float().permute(0, 2, 3, 1)) / (D ** 0.5)
    if causal:
        s = s.masked_fill(torch.arange(S, device=q.device)[None, :] > rows[:, None], float("-inf"))
    return torch.matmul(torch.softmax(s, dim=-1), v.float().permute(0, 2, 1, 3)).permute(0, 2, 1, 3)


def main():
    ref = load_ref()
    try:
        from flash_attn import flash_attn_func as fa
    except Exception:  # noqa: BLE001
        fa = None
    g = torch.Generator(device="cuda").manual_seed(0)
    H, D = 8, 128
    for S in (256, 512, 1024, 2048, 4096, 8192, 16384):
        q, k, v = (torch.randn(1, S, H, D, device="cuda", dtype=torch.bfloat16, generator=g) for _ in range(3))
        rows = torch.arange(S, device="cuda") if S <= 2048 else torch.unique(torch.cat([
            torch.arange(0, 256, device="cuda"), torch.randint(0, S, (1024,), device="cuda", generator=g), torch.tensor([S - 1], device="cuda")]))
        for causal in (False, True):
            o32 = rows_oracle(q, k, v, rows, causal)
            rec = {"S": S, "causal": causal, "rows_checked": int(rows.numel()), "tol": 1e-2}
            for name, fn in (("ours", lambda: mfa.flash_attn_func(q, k, v, causal=causal)),
                             ("reference", (lambda: ref.mini_flash_attention_forward(q, k, v, None, causal, -1, -1)) if ref else None),
                             ("flash_attn", (lambda: fa(q, k, v, causal=causal)) if fa else None)):
                if fn is None:
                    continue
                d = (fn()[:, rows].float() - o32).abs()
                big = d > 1e-2
                rec[name] = {"max_abs": round(float(d.max()), 5), "mean_abs": round(float(d.mean()), 7),
                             "frac_of_tol": round(float(d.max()) / 1e-2, 3), "elements_over_tol": int(big.sum()),
                             "max_abs_where_|o|<=1": round(float(d[o32.abs() <= 1].max()), 5)}
            print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
