"""`dllama inference | chat | worker` (reference: src/dllama.cpp).

``inference`` keeps the reference's benchmark output format: one ``🔷️ Eval`` line per prompt chunk,
one ``🔶 Pred`` line per generated token, and the final Evaluation / Prediction tokens/s block
(src/dllama.cpp:57-113).  Timings are CUDA-event device times of each forward; "Sync" is reported
as 0 because synchronisation is fused into the kernels (there is no separate sync step to time);
Sent/Recv kB count the peer-memory bytes this rank pushes/receives per forward.
"""
from __future__ import annotations

import sys
import time
from typing import List

import numpy as np

from ..parallel.control import OP_FORWARD
from ..text import native as T
from .app import USAGE, AppArgs, InferenceContext, run_inference_app, run_worker_app


class RootInference:
    """Root-side forward: publish the batch descriptor to the workers, then run the local engine
    (reference: RootLlmInference, src/app.cpp:169-209)."""

    def __init__(self, ctx: InferenceContext):
        self.ctx = ctx
        self.engine = ctx.engine
        self.slot = self.engine.alloc_slot()
        self.is_cuda = self.engine.device.type == "cuda"
        self.last_ms = 0.0

    def forward(self, tokens: List[int], pos: int, want_logits: bool = True):
        n = len(tokens)
        positions = list(range(pos, pos + n))
        slots = [self.slot] * n
        lf = n - 1
        if self.ctx.control is not None:
            self.ctx.control.send(OP_FORWARD, tokens, positions, slots, lf, False, [])
        if self.is_cuda:
            import torch
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            logits = self.engine.forward(tokens, positions, slots, logits_from=lf)
            e1.record()
            out = logits[0].float().cpu().numpy() if want_logits else None
            torch.cuda.synchronize()
            self.last_ms = e0.elapsed_time(e1)
        else:
            t0 = time.perf_counter()
            logits = self.engine.forward(tokens, positions, slots, logits_from=lf)
            out = logits[0].float().numpy().copy() if want_logits else None
            self.last_ms = (time.perf_counter() - t0) * 1e3
        return out

    def net(self):
        """(sync ms, sent bytes, received bytes) of the last forward on the TCP-synchronised CPU backend, else None."""
        f = getattr(self.engine, "network_stats", None)
        if f is None:
            return None
        st = f(reset=True)
        return st["sync_us"] / 1e3, st["sent_bytes"], st["recv_bytes"]

    def sync_bytes_per_row(self) -> int:
        """Peer-memory bytes this rank pushes per activation row per forward (partial sums + logits)."""
        e, h = self.engine, self.engine.h
        N = getattr(e, "world", 1)
        if N <= 1:
            return 0
        return (N - 1) * (2 * h.n_layers * h.dim * 4 + (h.vocab_size // N) * 4)


def inference(ctx: InferenceContext):
    a = ctx.args
    if a.prompt is None:
        raise ValueError("Prompt is required")
    if a.steps == 0:
        raise ValueError("Number of steps is required")
    tok = ctx.tokenizer
    input_tokens = tok.encode(a.prompt, add_bos=True, add_special_tokens=False)
    n_in = len(input_tokens)
    if n_in > ctx.header.seq_len:
        raise ValueError("The number of prompt tokens is greater than the sequence length")
    if n_in > a.steps:
        raise ValueError("The number of prompt tokens is greater than the number of steps")
    inf = RootInference(ctx)
    chunk = min(a.n_batches, ctx.engine.max_rows)
    sent = recv = 0
    per_row = inf.sync_bytes_per_row()
    eval_ms = pred_ms = 0.0
    print(a.prompt)
    pos = 0
    logits = None
    while pos < n_in - 1:                                  # evaluate all but the last prompt token
        n = min(chunk, n_in - 1 - pos)
        inf.forward(input_tokens[pos:pos + n], pos, want_logits=False)
        pos += n
        net = inf.net()
        sync_ms = 0.0
        if net is not None:                                # measured on the sockets (CPU backend)
            sync_ms, ds, dr = net
            sent += ds
            recv += dr
        else:
            sent += per_row * n
            recv += per_row * n
        print("🔷️ Eval%5d ms Sync%5d ms | Sent%6d kB Recv%6d kB | (%d tokens)" % (
            round(inf.last_ms), round(sync_ms), sent // 1024, recv // 1024, n), flush=True)
        eval_ms += inf.last_ms
    token = input_tokens[pos]                              # the last prompt token feeds the first prediction
    tok.reset_decoder()
    max_pos = min(ctx.header.seq_len, a.steps)
    n_pred = 0
    while pos < max_pos:
        logits = inf.forward([token], pos)
        token = ctx.sampler.sample(logits)
        piece = tok.decode(token)
        net = inf.net()
        sync_ms = 0.0
        if net is not None:
            sync_ms, ds, dr = net
            sent += ds
            recv += dr
        else:
            sent += per_row
            recv += per_row
        print("🔶 Pred%5d ms Sync%5d ms | Sent%6d kB Recv%6d kB | %s" % (
            round(inf.last_ms), round(sync_ms), sent // 1024, recv // 1024, piece if piece is not None else "~"), flush=True)
        pred_ms += inf.last_ms
        pos += 1
        n_pred += 1
    n_eval = n_in - 1
    print()
    print("Evaluation")
    print("   nBatches: %d" % chunk)
    print("    nTokens: %d" % n_eval)
    if n_eval > 0 and eval_ms > 0:
        print("   tokens/s: %3.2f (%3.2f ms/tok)" % (n_eval * 1000 / eval_ms, eval_ms / n_eval))
    print("Prediction")
    print("    nTokens: %d" % n_pred)
    if n_pred > 0 and pred_ms > 0:
        print("   tokens/s: %3.2f (%3.2f ms/tok)" % (n_pred * 1000 / pred_ms, pred_ms / n_pred))


def chat(ctx: InferenceContext):
    a, tok = ctx.args, ctx.tokenizer
    seq_len = ctx.header.seq_len
    stops = [tok.piece(t).decode("utf-8", "replace") for t in tok.eos_token_ids]
    gen = T.ChatTemplateGenerator(a.chat_template_type, tok.chat_template, stops[0] if stops else "")
    for s in stops:
        print(f"🛑 Stop: {s}")
    max_stop = max([len(s) for s in stops] + [1])
    eos = T.EosDetector(tok.eos_token_ids, stops, max_stop, max_stop)
    inf = RootInference(ctx)
    chunk = min(a.n_batches, ctx.engine.max_rows)

    def read(prompt: str) -> str:
        sys.stdout.write(prompt)
        sys.stdout.flush()
        line = sys.stdin.readline()
        if line == "":
            raise EOFError
        return line.rstrip("\n")

    items = []
    try:
        sys_prompt = read("💻 System prompt (optional): ")
    except EOFError:
        return
    if sys_prompt:
        items.append(("system", sys_prompt))
    pos = 0
    while pos < seq_len:
        try:
            user = ""
            while not user:
                user = read("\n👱 User\n> ")
        except EOFError:
            return
        items.append(("user", user))
        content, public = gen.generate(items, True)
        tokens = tok.encode(content, add_bos=(pos == 0), add_special_tokens=True)
        end = min(seq_len, pos + len(tokens) - 1)
        i = 0
        while pos < end:                                   # prefill all but the last prompt token
            n = min(chunk, end - pos)
            inf.forward(tokens[i:i + n], pos, want_logits=False)
            i += n
            pos += n
        token = tokens[i] if i < len(tokens) else tokens[-1]
        tok.reset_decoder()
        print("\n🤖 Assistant")
        if public:
            sys.stdout.write(public)
        while pos < seq_len:
            logits = inf.forward([token], pos)
            token = ctx.sampler.sample(logits)
            piece = tok.decode(token)
            kind = eos.append(token, piece)
            if kind in (T.EOS_NO, T.EOS_YES):
                delta = eos.get_delta()
                if delta:
                    sys.stdout.write(delta)
                    sys.stdout.flush()
                eos.reset()
            pos += 1
            if kind == T.EOS_YES:
                break
        items = []
    print("(end of context)")


def main(argv=None) -> int:
    argv = list(sys.argv if argv is None else argv)
    try:
        args = AppArgs.parse(argv, True)
        if args.help or args.mode is None:
            print(USAGE)
            return 0
        if args.mode == "inference":
            args.benchmark = True
            run_inference_app(args, inference)
        elif args.mode == "chat":
            run_inference_app(args, chat)
        elif args.mode == "worker":
            run_worker_app(args)
        else:
            raise ValueError("Unsupported mode")
    except (ValueError, RuntimeError, FileNotFoundError) as e:
        print(f"🚨 Critical error: {e}")
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
