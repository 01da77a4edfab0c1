"""SSP staleness service with the reference's API — ``SspManager(num_worker, max_stale)``,
``create_rpc_server(host)`` / ``create_rpc_client(host)``, ``client.check_staleness(task_index,
local_step) -> 1`` (inception/ssp_manager.py:14-58, Thrift IDL in inception/ssp/CheckStaleness.py:21).

The reference runs a Thrift ``TNonblockingServer`` (framed binary protocol, port 8000, one pool thread per
worker) on PS-0's host; the handler records the caller's clock and sleeps in 100 ms polls while
``local_step > min(clocks) + max_stale``.  Here the wire format is a fixed 8-byte request (two big-endian
int32) and a 4-byte reply over plain TCP, served by one thread per connection; blocked callers are woken by a
condition variable the moment the slowest worker reports (no 100 ms granularity) and the state is
lock-protected (the reference mutates it from a thread pool without a lock).

This service is used by the CPU / multi-host plumbing mode.  On one NVSwitch box the same protocol runs on
clock words in peer-mapped HBM (``dtf_ssp_clock`` in csrc/ps_update.cu, ``SymmComm.ssp_wait``).
"""
from __future__ import annotations

import socket
import socketserver
import struct
import threading
import time
from typing import List, Optional

from ..config import SSP_PORT
from ..utils.logs import log


class SspHandlerState:
    """Clock table + the blocking rule (reference SspHandler.check_staleness, ssp_manager.py:21-29)."""

    def __init__(self, num_worker: int, max_stale: int):
        self.local_step_list: List[int] = [0] * num_worker
        self.max_stale = max_stale
        self.cv = threading.Condition()
        self.max_lead_seen = 0
        self.closed = False

    def check_staleness(self, task_index: int, local_step: int, timeout: Optional[float] = None) -> int:
        t0 = time.time()
        with self.cv:
            self.local_step_list[task_index] = local_step
            self.cv.notify_all()
            warned = False
            while not self.closed and local_step > min(self.local_step_list) + self.max_stale:
                if not warned:
                    log("------ Hey! You are too fast! Wait! task %d step %d min %d ------"
                        % (task_index, local_step, min(self.local_step_list)))
                    warned = True
                if timeout is not None and time.time() - t0 > timeout:
                    return 0
                self.cv.wait(0.1)
            self.max_lead_seen = max(self.max_lead_seen, local_step - min(self.local_step_list))
        return 1

    def close(self):
        with self.cv:
            self.closed = True
            self.cv.notify_all()


class _Conn(socketserver.BaseRequestHandler):
    def handle(self):
        st: SspHandlerState = self.server.state  # type: ignore[attr-defined]
        sock: socket.socket = self.request
        sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        while True:
            buf = b""
            while len(buf) < 8:
                chunk = sock.recv(8 - len(buf))
                if not chunk:
                    return
                buf += chunk
            task, step = struct.unpack(">ii", buf)
            if task < 0:                      # control message: shut the server down
                st.close()
                threading.Thread(target=self.server.shutdown, daemon=True).start()
                sock.sendall(struct.pack(">i", 1))
                return
            sock.sendall(struct.pack(">i", st.check_staleness(task, step)))


class _Server(socketserver.ThreadingTCPServer):
    allow_reuse_address = True
    daemon_threads = True


class SspRpcServer:
    def __init__(self, host: str, port: int, state: SspHandlerState):
        self.server = _Server((host, port), _Conn)
        self.server.state = state  # type: ignore[attr-defined]
        self.state = state
        self.port = self.server.server_address[1]

    def serve(self) -> None:
        """Blocks (the reference calls this on PS task 0 before ``server.join()``)."""
        self.server.serve_forever(poll_interval=0.05)

    def serve_in_background(self) -> threading.Thread:
        t = threading.Thread(target=self.serve, daemon=True)
        t.start()
        return t

    def stop(self) -> None:
        self.state.close()
        self.server.shutdown()
        self.server.server_close()


class SspRpcClient:
    def __init__(self, host: str, port: int, connect_timeout: float = 30.0):
        t0 = time.time()
        while True:
            try:
                self.sock = socket.create_connection((host, port), timeout=5.0)
                break
            except OSError:
                if time.time() - t0 > connect_timeout:
                    raise
                time.sleep(0.1)
        self.sock.settimeout(None)
        self.sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)

    def check_staleness(self, task_index: int, local_step: int) -> int:
        self.sock.sendall(struct.pack(">ii", task_index, local_step))
        buf = b""
        while len(buf) < 4:
            chunk = self.sock.recv(4 - len(buf))
            if not chunk:
                raise ConnectionError("SSP server closed the connection")
            buf += chunk
        return struct.unpack(">i", buf)[0]

    def shutdown_server(self) -> None:
        try:
            self.check_staleness(-1, 0)
        except Exception:
            pass

    def close(self) -> None:
        try:
            self.sock.close()
        except OSError:
            pass


class SspManager:
    def __init__(self, num_worker: int, max_stale: int, port: int = SSP_PORT):
        self.num_worker = num_worker
        self.max_stale = max_stale
        self.port = port
        self.state = SspHandlerState(num_worker, max_stale)

    def create_rpc_server(self, server_host: str) -> SspRpcServer:
        return SspRpcServer(server_host, self.port, self.state)

    def create_rpc_client(self, server_host: str) -> SspRpcClient:
        return SspRpcClient(server_host, self.port)


def main(argv=None) -> int:
    """Manual probe of a running SSP service — the role of the reference's generated command-line client
    (``inception/ssp/CheckStaleness-remote``: ``[-h host[:port]] check_staleness <task_index> <local_step>``), or a
    stand-alone server for experiments:

        python -m distributed_tensorflow_models_b200.parallel.ssp_manager -h 10.0.0.1:8000 check_staleness 2 17
        python -m distributed_tensorflow_models_b200.parallel.ssp_manager serve --workers 4 --max_stale 5 [--port 8000]
    """
    import argparse
    import sys
    ap = argparse.ArgumentParser(prog="ssp_manager", add_help=False, description=main.__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--help", action="help")
    ap.add_argument("-h", dest="host", default="localhost:%d" % SSP_PORT, help="host[:port] of the service")
    sub = ap.add_subparsers(dest="cmd", required=True)
    c = sub.add_parser("check_staleness")
    c.add_argument("task_index", type=int)
    c.add_argument("local_step", type=int)
    s = sub.add_parser("serve")
    s.add_argument("--workers", type=int, required=True)
    s.add_argument("--max_stale", type=int, default=5)
    s.add_argument("--port", type=int, default=SSP_PORT)
    args = ap.parse_args(sys.argv[1:] if argv is None else argv)
    if args.cmd == "serve":
        SspManager(args.workers, args.max_stale, args.port).create_rpc_server("0.0.0.0").serve()
        return 0
    host, _, port = args.host.partition(":")
    client = SspRpcClient("127.0.0.1" if host == "localhost" else host, int(port or SSP_PORT), connect_timeout=5.0)
    try:
        print(client.check_staleness(args.task_index, args.local_step))
    finally:
        client.close()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
