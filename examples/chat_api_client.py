#!/usr/bin/env python3
"""Minimal client of dllama-api (the counterpart of the reference's examples/chat-api-client.js):
asks two questions, prints the answers and the token usage; `--stream` shows SSE streaming and
`--users N` fires N concurrent requests to show continuous batching.

    HOST=127.0.0.1 PORT=9990 python examples/chat_api_client.py [--stream] [--users 8]
"""
import json
import os
import sys
import threading
import time
import urllib.request

HOST, PORT = os.environ.get("HOST", "127.0.0.1"), int(os.environ.get("PORT", "9990"))
URL = f"http://{HOST}:{PORT}/v1/chat/completions"


def chat(messages, max_tokens=64, stream=False):
    body = json.dumps({"messages": messages, "temperature": 0.7, "stop": ["<|eot_id|>"], "max_tokens": max_tokens,
                       "stream": stream}).encode()
    req = urllib.request.Request(URL, data=body, headers={"Content-Type": "application/json"})
    with urllib.request.urlopen(req) as r:
        if not stream:
            return json.load(r)
        text = ""
        for raw in r:
            line = raw.decode("utf-8").strip()
            if line.startswith("data:") and line != "data: [DONE]":
                delta = json.loads(line[5:])["choices"][0]["delta"].get("content", "")
                sys.stdout.write(delta)
                sys.stdout.flush()
                text += delta
        print()
        return {"choices": [{"message": {"content": text}}], "usage": None}


def ask(system, user, stream):
    print(f"> system: {system}\n> user: {user}")
    res = chat([{"role": "system", "content": system}, {"role": "user", "content": user}], stream=stream)
    if not stream:
        print(res["choices"][0]["message"]["content"])
    print(res.get("usage"), "\n")


def main():
    stream = "--stream" in sys.argv
    if "--users" in sys.argv:
        n = int(sys.argv[sys.argv.index("--users") + 1])
        out, t0 = [None] * n, time.time()

        def one(i):
            out[i] = chat([{"role": "user", "content": f"Count to {i + 3}."}], max_tokens=48)
        ts = [threading.Thread(target=one, args=(i,)) for i in range(n)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        toks = sum(o["usage"]["completion_tokens"] for o in out)
        print(f"{n} concurrent users: {toks} tokens in {time.time() - t0:.2f}s")
        return
    ask("You are an excellent math teacher.", "What is 1 + 2?", stream)
    ask("You are a romantic.", "Where is Europe?", stream)


if __name__ == "__main__":
    main()
