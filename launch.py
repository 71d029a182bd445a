#!/usr/bin/env python3
"""Model launcher: download a converted model + tokenizer from the model zoo, write a run script and
optionally start `dllama chat` (reference: launch.py -- zoo :16-47, multi-part download with retries
:53-87, run script + start :121-161).

    python launch.py <model> [--gpus N] [--api] [--yes]
    python launch.py --list

The zoo points at the same published `.m` / `.t` artefacts as the reference (they are bit-compatible
with this runtime).  Downloads need network access; on an air-gapped box use
`python -m dllama_b200.formats.mfile`-style random models (see bench.py) instead.
"""
from __future__ import annotations

import os
import sys
import time
from urllib.request import urlopen

HF = "https://huggingface.co/b4rtaz"


def _parts(n):
    return [chr(97 + i // 26) + chr(97 + i % 26) for i in range(n)]


def _one(repo, model, tok):
    return ([f"{HF}/{repo}/resolve/main/{model}?download=true"], f"{HF}/{repo}/resolve/main/{tok}?download=true")


def _many(repo, prefix, n, tok):
    return ([f"{HF}/{repo}/resolve/main/{prefix}{s}?download=true" for s in _parts(n)],
            f"{HF}/{repo}/resolve/main/{tok}?download=true")


# name -> (model part urls, tokenizer url, weights type, buffer type, extra args)
MODELS = {
    "llama3_1_8b_instruct_q40": (*_one("Llama-3_1-8B-Q40-Instruct-Distributed-Llama", "dllama_model_llama3.1_instruct_q40.m",
                                      "dllama_tokenizer_llama_3_1.t"), "q40", "q80", "--max-seq-len 4096"),
    "llama3_1_405b_instruct_q40": (*_many("Llama-3_1-405B-Q40-Instruct-Distributed-Llama", "dllama_model_llama31_405b_q40_", 56,
                                         "dllama_tokenizer_llama_3_1.t"), "q40", "q80", "--max-seq-len 4096"),
    "llama3_2_1b_instruct_q40": (*_one("Llama-3_2-1B-Q40-Instruct-Distributed-Llama", "dllama_model_llama3.2-1b-instruct_q40.m",
                                      "dllama_tokenizer_llama3_2.t"), "q40", "q80", "--max-seq-len 4096"),
    "llama3_2_3b_instruct_q40": (*_one("Llama-3_2-3B-Q40-Instruct-Distributed-Llama", "dllama_model_llama3.2-3b-instruct_q40.m",
                                      "dllama_tokenizer_llama3_2.t"), "q40", "q80", "--max-seq-len 4096"),
    "llama3_3_70b_instruct_q40": (*_many("Llama-3_3-70B-Q40-Instruct-Distributed-Llama", "dllama_model_llama-3.3-70b_q40", 11,
                                        "dllama_tokenizer_llama-3.3-70b.t"), "q40", "q80", "--max-seq-len 4096"),
    "deepseek_r1_distill_llama_8b_q40": (*_one("DeepSeek-R1-Distill-Llama-8B-Distributed-Llama",
                                              "dllama_model_deepseek-r1-distill-llama-8b_q40.m",
                                              "dllama_tokenizer_deepseek-r1-distill-llama-8b.t"), "q40", "q80",
                                         "--max-seq-len 4096"),
}


def confirm(msg: str, assume_yes: bool) -> bool:
    if assume_yes:
        return True
    return input(f'{msg} ("Y" if yes): ').strip().upper() in ("Y", "YES")


def download(urls, path: str, assume_yes: bool, retries: int = 8):
    if os.path.isfile(path) and not confirm(f"{os.path.basename(path)} already exists, download again?", assume_yes):
        return
    done_mb = 0
    with open(path, "wb") as out:
        for url in urls:
            for attempt in range(retries):
                print(f"📄 {url}" + (f" (attempt {attempt + 1})" if attempt else ""))
                try:
                    with urlopen(url, timeout=60) as r:
                        while True:
                            chunk = r.read(1 << 20)
                            if not chunk:
                                break
                            out.write(chunk)
                            done_mb += 1
                            if done_mb % 64 == 0:
                                sys.stdout.write(f"\rDownloaded {done_mb} MB")
                                sys.stdout.flush()
                    break
                except Exception as e:                      # partial part: rewind and retry
                    print(f"\n🚨 download failed: {e}")
                    if attempt == retries - 1:
                        raise
                    time.sleep(2 * (attempt + 1))
    print(" ✅")


def main(argv):
    if len(argv) < 2 or argv[1] in ("-h", "--help"):
        print(__doc__)
        return 1
    if argv[1] == "--list":
        for name in MODELS:
            print(name)
        return 0
    name = argv[1].replace("-", "_")
    if name not in MODELS:
        print(f"Model is not supported: {name}\nAvailable: " + ", ".join(MODELS))
        return 1
    gpus = int(argv[argv.index("--gpus") + 1]) if "--gpus" in argv else 1
    api = "--api" in argv
    yes = "--yes" in argv
    model_urls, tok_url, wt, bt, extra = MODELS[name]
    root = os.path.dirname(os.path.abspath(__file__))
    folder = os.path.join(root, "models", name)
    os.makedirs(folder, exist_ok=True)
    model_path = os.path.join(folder, f"dllama_model_{name}.m")
    tok_path = os.path.join(folder, f"dllama_tokenizer_{name}.t")
    download(model_urls, model_path, yes)
    download([tok_url], tok_path, yes)
    exe = "dllama-api" if api else "dllama chat"
    launcher = f"python -m torch.distributed.run --nnodes=1 --nproc-per-node {gpus} --master-addr 127.0.0.1 " if gpus > 1 else "./"
    cmd = f"{launcher}{exe} --model {model_path} --tokenizer {tok_path} --buffer-float-type {bt} {extra}"
    script = os.path.join(root, f"run_{name}.sh")
    with open(script, "w") as f:
        f.write("#!/bin/sh\n\n" + cmd + "\n")
    os.chmod(script, 0o755)
    print(f"To run the model: sh {os.path.basename(script)}\n   {cmd}")
    if confirm("Do you want to run it now?", yes):
        if not os.path.exists(os.path.join(root, "distributed-llama-multiusers_b200", "lib", "libdllama_b200_cuda.so")):
            os.system(f"{sys.executable} {os.path.join(root, 'build.py')}")
        os.system(cmd)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
