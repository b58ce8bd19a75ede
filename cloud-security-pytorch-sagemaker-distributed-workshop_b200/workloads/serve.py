"""A stand-alone endpoint for the serving hooks: ``GET /ping`` and ``POST /invocations``.

The reference deploys its ``inference.py`` into SageMaker's PyTorch serving container
(``notebooks/1_pytorch_dist_native_cpu.ipynb:153-211``: ``PyTorchModel(entry_point='inference.py').deploy(...)`` then
``predictor.predict(...)``), which is what turns ``model_fn`` / ``input_fn`` / ``predict_fn`` / ``output_fn`` into an HTTP
endpoint.  Outside SageMaker there is no such container; this module is the same contract on the standard library's HTTP
server, so a ``model.pth`` written by the trainer can be served and queried the way the notebook does:

    python -m b200ddp.workloads.serve --model-dir /opt/ml/model --port 8080
    curl -s localhost:8080/ping
    curl -s -H 'Content-Type: application/json' -d '[[...3x32x32...]]' localhost:8080/invocations

``Content-Type`` selects ``input_fn``'s decoder (``application/x-npy`` or ``application/json``), ``Accept`` selects
``output_fn``'s encoder.  One model per process, loaded once; requests are served one at a time (the hooks are not
re-entrant by contract).
"""
from __future__ import annotations

import argparse
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

from . import inference


def make_server(model_dir: str, host: str = "127.0.0.1", port: int = 8080, hooks=inference) -> ThreadingHTTPServer:
    """Load the model with ``hooks.model_fn`` and return a server (not yet serving) bound to ``host:port`` (port 0 = any)."""
    model = hooks.model_fn(model_dir)
    lock = threading.Lock()

    class Handler(BaseHTTPRequestHandler):
        protocol_version = "HTTP/1.1"

        def _reply(self, code: int, body: bytes, ctype: str = "text/plain"):
            self.send_response(code)
            self.send_header("Content-Type", ctype)
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

        def do_GET(self):                                   # noqa: N802 (http.server's naming)
            if self.path.rstrip("/") == "/ping":
                self._reply(200, b"")                       # the container contract: 200 and an empty body when healthy
            else:
                self._reply(404, b"not found")

        def do_POST(self):                                  # noqa: N802
            if self.path.rstrip("/") != "/invocations":
                self._reply(404, b"not found")
                return
            n = int(self.headers.get("Content-Length", "0"))
            body = self.rfile.read(n)
            ctype = (self.headers.get("Content-Type") or "application/x-npy").split(";")[0].strip()
            accept = (self.headers.get("Accept") or "application/json").split(",")[0].split(";")[0].strip()
            if accept in ("*/*", ""):
                accept = "application/json"
            try:
                data = hooks.input_fn(body, ctype)
            except Exception as e:                          # undecodable request: the client's fault
                self._reply(415 if isinstance(e, ValueError) and "unsupported content type" in str(e) else 400,
                            ("bad request: %s" % e).encode())
                return
            try:
                with lock:
                    out = hooks.output_fn(hooks.predict_fn(data, model), accept)
            except Exception as e:
                self._reply(500, ("prediction failed: %s: %s" % (type(e).__name__, e)).encode())
                return
            payload = out if isinstance(out, (bytes, bytearray)) else str(out).encode()
            self._reply(200, bytes(payload), accept if accept == "application/x-npy" else "application/json")

        def log_message(self, fmt, *args):                  # quiet: one line per request would drown the trainer-style logs
            return

    return ThreadingHTTPServer((host, port), Handler)


def main(argv=None):
    ap = argparse.ArgumentParser(description="serve a trained model.pth: GET /ping, POST /invocations")
    ap.add_argument("--model-dir", required=True)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=8080)
    a = ap.parse_args(argv)
    srv = make_server(a.model_dir, a.host, a.port)
    print("serving %s on http://%s:%d (GET /ping, POST /invocations)" % (a.model_dir, *srv.server_address), flush=True)
    try:
        srv.serve_forever()
    except KeyboardInterrupt:
        pass
    finally:
        srv.server_close()


if __name__ == "__main__":
    main()
