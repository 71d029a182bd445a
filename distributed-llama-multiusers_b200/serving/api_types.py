"""OpenAI-style DTOs of the HTTP API (reference: src/api-types.hpp:10-177).  Plain dict builders."""
from __future__ import annotations

import time
import uuid
from dataclasses import dataclass, field
from typing import Dict, List, Optional


@dataclass
class InferenceParams:
    """(reference: struct InferenceParams, src/api-types.hpp:103-111)"""
    messages: List[Dict[str, str]]
    max_tokens: int = -1
    temperature: float = 0.8
    top_p: float = 0.9
    stop: List[str] = field(default_factory=list)
    stream: bool = False
    seed: int = 0
    include_usage: bool = False      # stream_options.include_usage: usage in the last SSE chunk (OpenAI)


def parse_chat_messages(items) -> List[Dict[str, str]]:
    if not isinstance(items, list):
        raise ValueError("`messages` must be a list")
    out = []
    for it in items:
        if not isinstance(it, dict) or "role" not in it or "content" not in it:
            raise ValueError("every message needs `role` and `content`")
        content = it["content"]
        if isinstance(content, list):           # OpenAI content-part arrays: keep the text parts
            content = "".join(p.get("text", "") for p in content if isinstance(p, dict))
        out.append({"role": str(it["role"]), "content": str(content)})
    return out


def parse_inference_params(body: dict, defaults: InferenceParams) -> InferenceParams:
    """(reference: ApiServer::parseRequest, src/dllama-api.cpp:291-313) -- here every field takes effect."""
    p = InferenceParams(messages=parse_chat_messages(body.get("messages")), max_tokens=defaults.max_tokens,
                        temperature=defaults.temperature, top_p=defaults.top_p, seed=defaults.seed)
    if "stream" in body:
        p.stream = bool(body["stream"])
    if body.get("temperature") is not None:
        p.temperature = float(body["temperature"])
    if body.get("top_p") is not None:
        p.top_p = float(body["top_p"])
    if body.get("seed") is not None:
        p.seed = int(body["seed"])
    if body.get("max_tokens") is not None:
        p.max_tokens = int(body["max_tokens"])
    elif body.get("max_completion_tokens") is not None:
        p.max_tokens = int(body["max_completion_tokens"])
    so = body.get("stream_options")
    if isinstance(so, dict) and so.get("include_usage"):
        p.include_usage = True
    stop = body.get("stop")
    if isinstance(stop, str):
        p.stop = [stop]
    elif isinstance(stop, list):
        p.stop = [str(s) for s in stop]
    return p


def new_completion_id() -> str:
    return "cmpl-" + uuid.uuid4().hex[:24]


def chat_usage(prompt_tokens: int, completion_tokens: int) -> dict:
    return {"prompt_tokens": prompt_tokens, "completion_tokens": completion_tokens,
            "total_tokens": prompt_tokens + completion_tokens}


def chat_completion(cid: str, model: str, content: str, finish_reason: str, usage: dict) -> dict:
    return {"id": cid, "object": "chat.completion", "created": int(time.time()), "model": model, "usage": usage,
            "choices": [{"index": 0, "message": {"role": "assistant", "content": content},
                         "finish_reason": finish_reason}],
            # kept for the fork's web-ui, which reads `generated_text` (web-ui/app.js:40)
            "generated_text": content}


def chat_completion_chunk(cid: str, model: str, delta: Optional[dict], finish_reason: Optional[str]) -> dict:
    return {"id": cid, "object": "chat.completion.chunk", "created": int(time.time()), "model": model,
            "choices": [{"index": 0, "delta": delta or {}, "finish_reason": finish_reason}]}


def model_list(model_name: str) -> dict:
    return {"object": "list",
            "data": [{"id": model_name, "object": "model", "created": 0, "owned_by": "user"}]}
