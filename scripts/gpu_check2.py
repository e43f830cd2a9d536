"""Development probe: attention + encoder parity on the GPU."""
import os, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import q3a_b200  # noqa
from q3a_b200 import binding as B, synth
from oracle import mel_oracle as MO, encoder_oracle as EO

B.load()
torch.manual_seed(0)

def attn_ref(qkv, H, cu):
    T = qkv.shape[0]; d = H * 64
    q, k, v = qkv[:, :d].float(), qkv[:, d:2*d].float(), qkv[:, 2*d:].float()
    out = torch.zeros(T, d, device=qkv.device)
    for s, e in zip(cu[:-1], cu[1:]):
        qs = q[s:e].view(e-s, H, 64).transpose(0, 1); ks = k[s:e].view(e-s, H, 64).transpose(0, 1); vs = v[s:e].view(e-s, H, 64).transpose(0, 1)
        p = torch.softmax(qs @ ks.transpose(1, 2) * 0.125, dim=-1)
        out[s:e] = (p @ vs).transpose(0, 1).reshape(e-s, d)
    return out

for opd, dt in ((B.OP_BF16, torch.bfloat16), (B.OP_F16, torch.float16)):
    for H, cu in ((2, [0, 128]), (2, [0, 16]), (3, [0, 100]), (14, [0, 390]), (4, [0, 104, 208, 250]), (2, [0, 1, 130, 400, 1196 + 400])):
        T = cu[-1]
        qkv = (torch.randn(T, 3 * H * 64, device="cuda") * 1.5).to(dt)
        out = torch.zeros(T, H * 64, device="cuda", dtype=dt)
        B.test_attention(qkv, out, H, cu, opd)
        torch.cuda.synchronize()
        want = attn_ref(qkv, H, cu)
        err = (out.float() - want).abs().max().item()
        print(f"attn {dt} H={H} cu={cu} maxabs={err:.3e} (ref max {want.abs().max().item():.2f})", flush=True)

def cmp(name, got, want):
    cos = float((got * want).sum() / (np.linalg.norm(got) * np.linalg.norm(want) + 1e-30))
    rel = float(np.abs(got - want).max() / (np.abs(want).max() + 1e-30))
    print(f"{name}: shape={got.shape} cos={cos:.6f} maxrel={rel:.3e}", flush=True)

tmp = tempfile.mkdtemp()
filt = MO.mel_filters_numpy()
# ---- tiny tower, all modes
shape = synth.tiny_shape()
w = synth.random_weights(shape, 7)
for ftype in ("f32", "f16", "q8_0"):
    path = synth.write_gguf(os.path.join(tmp, f"tiny_{ftype}.gguf"), shape, w, ftype)
    W = EO.Weights({k: v for k, v in synth.stored_weights(w, ftype).items() if "#" not in k}, ftype)
    for opd in (B.OP_F16, B.OP_BF16):
        enc = B.Encoder(path, 0, opd)
        for n in (16000 * 3 + 777, 16000 + 5, 100 * 160):
            pcm = synth.synth_audio(11, n)
            mel = MO.log_mel_numpy(pcm, filt)
            for mode, variant in ((B.MODE_ASR, "asr"), (B.MODE_ALIGNER, "aligner")):
                got = enc.encode_batch([mel], mode)[0]
                want = EO.encode(mel, W, n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads, variant=variant)
                cmp(f"tiny {ftype} op={opd} n={n} {variant}", got, want)
        enc.close()
# ---- real ASR tower, 5 s and 30 s
shape = synth.ASR_06B
t0 = time.time(); w = synth.random_weights(shape, 0); print("weights", time.time() - t0, flush=True)
path = synth.write_gguf(os.path.join(tmp, "asr_f16.gguf"), shape, w, "f16")
W = EO.Weights(synth.stored_weights(w, "f16"), "f16")
for opd in (B.OP_F16, B.OP_BF16):
    t0 = time.time(); enc = B.Encoder(path, 0, opd); print("load s", time.time() - t0, "bytes", enc.device_bytes(), flush=True)
    for n in (16000 * 5 + 123, 480000):
        pcm = synth.synth_audio(100, n)
        mel = MO.log_mel_numpy(pcm, filt)
        got = enc.encode_batch([mel], B.MODE_ASR)[0]
        t0 = time.time()
        want = EO.encode(mel, W, n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads, variant="asr")
        print("oracle s", time.time() - t0)
        cmp(f"asr0.6b f16 op={opd} n={n}", got, want)
        if n == 480000:
            ideal = EO.encode(mel, W, n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads, variant="asr", nm=EO.IDEAL)
            cmp("   oracle vs ideal-f64", want, ideal.astype(np.float32))
            cmp("   ours vs ideal-f64", got, ideal.astype(np.float32))
    # batch of 64 x 30 s through the pcm path, timing
    fe = B.Frontend(0)
    clips = [synth.synth_audio(200 + i, 480000) for i in range(64)]
    enc.set_profiling(True)
    for it in range(3):
        t0 = time.time(); outs = enc.encode_pcm_batch(fe, clips, B.MODE_ASR); dt = time.time() - t0
        tm = enc.timings()
        print(f"batch64 e2e {dt*1e3:.1f} ms -> {64*30/dt:.0f} audio-s/s | mel {tm.mel_ms:.2f} stem {tm.conv_chunk_ms:.2f} xfmr {tm.transformer_ms:.2f} gemm {tm.gemm_ms:.2f} ({tm.gemm_flops/tm.gemm_ms/1e9:.0f} TF/s) attn {tm.attention_ms:.2f} ({tm.attention_flops/max(tm.attention_ms,1e-9)/1e9:.0f} TF/s) other {tm.other_ms:.2f} launches {tm.total_launches}", flush=True)
    one = enc.encode_pcm_batch(fe, clips[:1], B.MODE_ASR)[0]
    print("batch invariance (clip 0 alone vs in batch): bit-identical =", np.array_equal(one, outs[0]), "maxdiff", np.abs(one - outs[0]).max())
    mel0 = MO.log_mel_numpy(clips[0], filt)
    viamel = enc.encode_batch([mel0], B.MODE_ASR)[0]
    cmp("pcm path vs mel path", outs[0], viamel)
    enc.close()
