"""Python mirror of the reference's caller, for the contract tests (test infrastructure, not part of the product).

`ai_suggest` behaves like the function of the same name at web/streamlit_app.py:89-101: same environment
variables (OLLAMA_URL, LLM_MODEL, web/streamlit_app.py:27-28), same prompt wrapper (:93), same request body
(:91-95), same three outcomes (the stripped `response` on HTTP 200, "(LLM error)" on any other status,
"(LLM unavailable: ...)" on an exception).  Pointing OLLAMA_URL at llmd is the whole integration.
"""
import os

import requests

PROMPT_WRAPPER = ("You are a helpful assistant. Draft a concise, friendly reply to the following message:\n\n"
                  "{message}\n\nReply:")


def suggestion_request(message, model=None):
    """The JSON body the reference posts (web/streamlit_app.py:92-94)."""
    return {"model": model or os.environ.get("LLM_MODEL", "llama3.1"),
            "prompt": PROMPT_WRAPPER.format(message=message), "stream": False}


def ai_suggest(prompt, timeout=60):
    base = os.environ.get("OLLAMA_URL", "http://127.0.0.1:11434")
    try:
        resp = requests.post(f"{base}/api/generate", json=suggestion_request(prompt), timeout=timeout)
    except Exception as e:  # connection refused, timeout, ...
        return f"(LLM unavailable: {e})"
    if resp.status_code != 200:
        return "(LLM error)"
    return resp.json().get("response", "").strip()
