// Browser chat client for dllama-api (OpenAI-style /v1/chat/completions).
// Keeps the whole conversation, supports SSE streaming, and -- for compatibility with the fork's
// original server -- falls back to a bare `generated_text` field when `choices` is absent.
(() => {
  const $ = (id) => document.getElementById(id);
  const log = $('log'), input = $('input'), form = $('form'), send = $('send');
  const history = [];

  function bubble(role, text) {
    const el = document.createElement('div');
    el.className = `msg ${role}`;
    el.textContent = text;
    log.appendChild(el);
    log.scrollTop = log.scrollHeight;
    return el;
  }

  async function readStream(resp, el) {
    const reader = resp.body.getReader();
    const dec = new TextDecoder();
    let buf = '', text = '';
    for (;;) {
      const { value, done } = await reader.read();
      if (done) break;
      buf += dec.decode(value, { stream: true });
      let cut;
      while ((cut = buf.indexOf('\n\n')) >= 0) {
        const event = buf.slice(0, cut).trim();
        buf = buf.slice(cut + 2);
        if (!event.startsWith('data:')) continue;
        const payload = event.slice(5).trim();
        if (payload === '[DONE]') return text;
        const delta = JSON.parse(payload).choices?.[0]?.delta?.content;
        if (delta) { text += delta; el.textContent = text; log.scrollTop = log.scrollHeight; }
      }
    }
    return text;
  }

  async function ask(prompt) {
    history.push({ role: 'user', content: prompt });
    bubble('user', prompt);
    const el = bubble('assistant', '…');
    const stream = $('stream').checked;
    const t0 = performance.now();
    const resp = await fetch(`${$('server').value.replace(/\/$/, '')}/v1/chat/completions`, {
      method: 'POST',
      headers: { 'Content-Type': 'application/json' },
      body: JSON.stringify({ messages: history, max_tokens: 256, stream }),
    });
    if (!resp.ok) throw new Error(`HTTP ${resp.status}: ${await resp.text()}`);
    let text, usage = null;
    if (stream) {
      text = await readStream(resp, el);
    } else {
      const data = await resp.json();
      text = data.choices?.[0]?.message?.content ?? data.generated_text ?? '';
      usage = data.usage;
      el.textContent = text;
    }
    history.push({ role: 'assistant', content: text });
    const meta = document.createElement('div');
    meta.className = 'meta';
    const secs = ((performance.now() - t0) / 1000).toFixed(2);
    meta.textContent = usage ? `${usage.completion_tokens} tokens in ${secs}s` : `${secs}s`;
    el.appendChild(meta);
  }

  form.addEventListener('submit', async (e) => {
    e.preventDefault();
    const prompt = input.value.trim();
    if (!prompt) return;
    input.value = '';
    send.disabled = true;
    try { await ask(prompt); } catch (err) { bubble('error', String(err)); history.pop(); } finally { send.disabled = false; input.focus(); }
  });
  input.addEventListener('keydown', (e) => {
    if (e.key === 'Enter' && !e.shiftKey) { e.preventDefault(); form.requestSubmit(); }
  });
})();
