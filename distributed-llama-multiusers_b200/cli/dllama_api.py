"""`dllama-api`: multi-user OpenAI-style HTTP server (reference: src/dllama-api.cpp:367-410)."""
from __future__ import annotations

import signal
import sys

from ..serving import api_types as A
from ..serving.api import ApiServer, ApiState
from ..serving.batch_engine import BatchEngine
from .app import USAGE, AppArgs, InferenceContext, run_inference_app


def server(ctx: InferenceContext):
    a = ctx.args
    defaults = A.InferenceParams(messages=[], temperature=a.temperature, top_p=a.topp, seed=a.seed)
    import os
    if a.api_backend == "native":
        # C++ front end (csrc/host/server.cpp): sockets, HTTP/JSON/SSE, tokenizer, scheduler, detokeniser in one
        # I/O thread; this thread only runs the forward passes
        from ..serving.native_server import NativeApiServer
        srv = NativeApiServer(ctx.engine, ctx.tokenizer, os.path.basename(a.model_path), defaults, a.chat_template_type,
                              control=ctx.control, host="0.0.0.0", port=a.port)
        srv.warmup()                     # capture the CUDA graphs of every batch shape before taking traffic
        signal.signal(signal.SIGTERM, lambda *_: srv._stop.set())
        try:
            srv.serve_forever()
        except KeyboardInterrupt:
            pass
        finally:
            srv.close()
        return
    batch = BatchEngine(ctx.engine, ctx.tokenizer, control=ctx.control)
    batch.warmup()                       # capture the CUDA graphs of every batch shape before taking traffic
    batch.start()
    state = ApiState(batch, ctx.tokenizer, os.path.basename(a.model_path), defaults, a.chat_template_type)
    srv = ApiServer(state, "0.0.0.0", a.port)
    # (shutdown() blocks until serve_forever returns, and the handler runs on the thread that is inside it)
    import threading
    signal.signal(signal.SIGTERM, lambda *_: threading.Thread(target=srv.httpd.shutdown, daemon=True).start())
    try:
        srv.serve_forever()
    except KeyboardInterrupt:
        pass
    finally:
        ctx.control = None if ctx.control is None else ctx.control
        batch.stop()


def main(argv=None) -> int:
    argv = list(sys.argv if argv is None else argv)
    try:
        args = AppArgs.parse(argv, False)
        if args.help:
            print(USAGE)
            return 0
        run_inference_app(args, server)
    except (ValueError, RuntimeError, FileNotFoundError) as e:
        print(f"🚨 Critical error: {e}")
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
