"""Host-side transports: in-process loopback hub and a TCP star.

The reference's transport is socket.io over websockets in a star topology
(`/root/reference/src/server/asynchronousSGD_server.ts:50-80`,
`/root/reference/src/client/abstract_client.ts:148-173`): the
server reacts to `connection`, unicasts with `socket.emit`, broadcasts with
`server.sockets.emit`; a client `emit`s with an application-level ack guarded
by a timeout.  The same surface is offered here by two back ends:

  * `LoopbackHub` — everything in one process, single-threaded and
    deterministic (messages are queued and delivered by `pump()`), used by the
    unit tests and by the CPU "2 workers" plumbing config;
  * `TcpServerEndpoint` / `TcpClientEndpoint` — length-prefixed frames over
    TCP sockets for server and workers in different processes.

Between GPU ranks none of this is used: `parallel/` moves gradients and
weights through peer-mapped device memory and the only host traffic is
bootstrap (handle exchange) through `torch.distributed`.
"""
from __future__ import annotations

import collections
import itertools
import socket
import struct
import threading
import time
from typing import Any, Callable, Deque, Dict, List, Optional, Tuple

from .messages import MAX_FRAME_BYTES, Events, WireFormatError, decode, encode

__all__ = ["LoopbackHub", "LoopbackPeer", "LoopbackClientEndpoint", "TcpServerEndpoint", "TcpClientEndpoint",
           "AckTimeout", "ConnectTimeout"]


class AckTimeout(TimeoutError):
    pass


class ConnectTimeout(TimeoutError):
    pass


# ==============================================================================================
# in-process loopback

class LoopbackPeer:
    """Server-side view of one connected client (the `socket` of socket.io)."""

    def __init__(self, hub: "LoopbackHub", client: "LoopbackClientEndpoint", pid: int):
        self.hub, self.client, self.id = hub, client, pid
        self.handlers: Dict[str, Callable[[Any, Callable[[Any], None]], None]] = {}
        self.connected = True

    def on(self, event: str, handler: Callable[[Any, Callable[[Any], None]], None]) -> None:
        self.handlers[event] = handler

    def emit(self, event: str, msg: Any) -> None:
        if self.connected:
            self.hub._enqueue(lambda: self.client._deliver(event, msg))


class LoopbackHub:
    """Server endpoint + message queue.  `emit`s enqueue; `pump()` delivers in FIFO order."""

    def __init__(self, drop: Optional[Callable[[str, Any], bool]] = None):
        self._queue: Deque[Callable[[], None]] = collections.deque()
        self._conn_handlers: List[Callable[[LoopbackPeer], None]] = []
        self.peers: List[LoopbackPeer] = []
        self._ids = itertools.count()
        self.drop = drop            # fault injection: drop(event, msg) -> True loses a client->server message
        self.closed = False

    # server API ------------------------------------------------------------------------------
    def on_connection(self, handler: Callable[[LoopbackPeer], None]) -> None:
        self._conn_handlers.append(handler)

    def broadcast(self, event: str, msg: Any) -> None:       # server.sockets.emit
        for p in list(self.peers):
            p.emit(event, msg)

    def close(self) -> None:
        self.closed = True
        for p in list(self.peers):
            p.client.close()

    # plumbing --------------------------------------------------------------------------------
    def _enqueue(self, fn: Callable[[], None]) -> None:
        self._queue.append(fn)

    def pump(self, until: Optional[Callable[[], bool]] = None, max_events: int = 10_000_000) -> int:
        n = 0
        while self._queue and n < max_events:
            if until is not None and until():
                break
            self._queue.popleft()()
            n += 1
        return n

    def client(self) -> "LoopbackClientEndpoint":
        return LoopbackClientEndpoint(self)

    def _connect(self, client: "LoopbackClientEndpoint") -> LoopbackPeer:
        if self.closed:
            raise ConnectTimeout("server closed")
        peer = LoopbackPeer(self, client, next(self._ids))
        self.peers.append(peer)
        for h in self._conn_handlers:
            self._enqueue(lambda h=h: h(peer))
        return peer

    def _disconnect(self, peer: LoopbackPeer) -> None:
        if peer in self.peers:
            self.peers.remove(peer)
            peer.connected = False
            h = peer.handlers.get(Events.Disconnect)
            if h:
                self._enqueue(lambda: h(None, lambda _=None: None))


class LoopbackClientEndpoint:
    def __init__(self, hub: LoopbackHub):
        self.hub = hub
        self.peer: Optional[LoopbackPeer] = None
        self.handlers: Dict[str, List[Callable[[Any], None]]] = collections.defaultdict(list)
        self.inbox: Dict[str, Deque[Any]] = collections.defaultdict(collections.deque)

    def connect(self, timeout: float = 10.0) -> None:
        self.peer = self.hub._connect(self)

    def on(self, event: str, handler: Callable[[Any], None]) -> None:
        self.handlers[event].append(handler)

    def _deliver(self, event: str, msg: Any) -> None:
        hs = self.handlers.get(event)
        if hs:
            for h in list(hs):
                h(msg)
        else:
            self.inbox[event].append(msg)

    def wait_event(self, event: str, timeout: float = 10.0) -> Any:
        """One-shot wait (`fromEvent`, /root/reference/src/client/utils.ts:5-19)."""
        self.hub.pump(until=lambda: bool(self.inbox[event]))
        if not self.inbox[event]:
            raise ConnectTimeout(f"no '{event}' within {timeout}s")
        return self.inbox[event].popleft()

    def emit(self, event: str, msg: Any, ack_timeout: float = 5.0) -> Any:
        """Send with ack; pumps the hub until the server acknowledged."""
        if self.peer is None or not self.peer.connected:
            raise AckTimeout("not connected")
        box: List[Any] = []
        if self.hub.drop is not None and self.hub.drop(event, msg):
            raise AckTimeout(f"no ack for '{event}' within {ack_timeout}s")
        peer = self.peer

        def run() -> None:
            h = peer.handlers.get(event)
            if h is not None:
                h(msg, lambda v=True: box.append(v))

        self.hub._enqueue(run)
        self.hub.pump(until=lambda: bool(box))
        if not box:
            raise AckTimeout(f"no ack for '{event}' within {ack_timeout}s")
        return box[0]

    def close(self) -> None:
        if self.peer is not None:
            self.hub._disconnect(self.peer)
            self.peer = None


# ==============================================================================================
# TCP star

_HDR = struct.Struct("<IIB")     # payload length, message id, kind (0 = event, 1 = ack)


def _send_frame(sock: socket.socket, lock: threading.Lock, mid: int, kind: int, obj: Any) -> None:
    payload = encode(obj)
    with lock:
        sock.sendall(_HDR.pack(len(payload), mid, kind) + payload)


def _recv_exact(sock: socket.socket, n: int) -> Optional[bytes]:
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(min(n - len(buf), 1 << 20))
        if not chunk:
            return None
        buf.extend(chunk)
    return bytes(buf)


def _recv_frame(sock: socket.socket) -> Optional[Tuple[int, int, Any]]:
    hdr = _recv_exact(sock, _HDR.size)
    if hdr is None:
        return None
    n, mid, kind = _HDR.unpack(hdr)
    if n > MAX_FRAME_BYTES or kind not in (0, 1):
        return None                               # not our protocol (or an absurd length): drop the connection
    payload = _recv_exact(sock, n)
    if payload is None:
        return None
    try:
        return mid, kind, decode(payload)
    except WireFormatError:
        return None


class _TcpPeer:
    def __init__(self, server: "TcpServerEndpoint", sock: socket.socket, pid: int):
        self.server, self.sock, self.id = server, sock, pid
        self.handlers: Dict[str, Callable] = {}
        self.wlock = threading.Lock()
        self.connected = True

    def on(self, event: str, handler: Callable) -> None:
        self.handlers[event] = handler

    def emit(self, event: str, msg: Any) -> None:
        if not self.connected:
            return
        try:
            _send_frame(self.sock, self.wlock, 0, 0, (event, msg))
        except OSError:
            self.connected = False


class TcpServerEndpoint:
    """Accepts worker connections; handlers run one at a time (JS-like single-threaded semantics)."""

    def __init__(self, host: str = "127.0.0.1", port: int = 0):
        self._lsock = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self._lsock.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self._lsock.bind((host, port))
        self._lsock.listen(64)
        self.host, self.port = self._lsock.getsockname()
        self.peers: List[_TcpPeer] = []
        self._conn_handlers: List[Callable] = []
        self._lock = threading.RLock()          # serialises all server-side handlers
        self._ids = itertools.count()
        self.closed = False
        self._thread = threading.Thread(target=self._accept_loop, daemon=True)
        self._thread.start()

    @property
    def url(self) -> str:
        return f"tcp://{self.host}:{self.port}"

    def on_connection(self, handler: Callable) -> None:
        self._conn_handlers.append(handler)

    def broadcast(self, event: str, msg: Any) -> None:
        for p in list(self.peers):
            p.emit(event, msg)

    def _accept_loop(self) -> None:
        while not self.closed:
            try:
                sock, _ = self._lsock.accept()
            except OSError:
                return
            sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            peer = _TcpPeer(self, sock, next(self._ids))
            with self._lock:
                self.peers.append(peer)
                for h in self._conn_handlers:
                    h(peer)
            threading.Thread(target=self._reader, args=(peer,), daemon=True).start()

    def _reader(self, peer: _TcpPeer) -> None:
        while True:
            try:
                frame = _recv_frame(peer.sock)
            except OSError:
                frame = None
            if frame is None:
                break
            mid, kind, obj = frame
            if kind != 0 or not (isinstance(obj, tuple) and len(obj) == 2 and isinstance(obj[0], str)):
                break                             # not an event frame: protocol violation, drop the peer
            event, msg = obj

            def ack(v: Any = True, mid=mid) -> None:
                try:
                    _send_frame(peer.sock, peer.wlock, mid, 1, v)
                except OSError:
                    pass

            with self._lock:
                h = peer.handlers.get(event)
                if h is not None:
                    h(msg, ack)
        peer.connected = False
        with self._lock:
            if peer in self.peers:
                self.peers.remove(peer)
            h = peer.handlers.get(Events.Disconnect)
            if h:
                h(None, lambda _=None: None)
        try:
            peer.sock.close()
        except OSError:
            pass

    def pump(self, *a, **k) -> int:      # API compatibility with LoopbackHub
        return 0

    def close(self) -> None:
        self.closed = True
        try:
            self._lsock.close()
        except OSError:
            pass
        for p in list(self.peers):
            try:
                p.sock.shutdown(socket.SHUT_RDWR)
            except OSError:
                pass


class TcpClientEndpoint:
    def __init__(self, url: str):
        assert url.startswith("tcp://"), url
        host, port = url[len("tcp://"):].rsplit(":", 1)
        self.addr = (host, int(port))
        self.sock: Optional[socket.socket] = None
        self.handlers: Dict[str, List[Callable]] = collections.defaultdict(list)
        self.inbox: Dict[str, Deque[Any]] = collections.defaultdict(collections.deque)
        self._cv = threading.Condition()
        self._acks: Dict[int, Any] = {}
        self._ids = itertools.count(1)
        self._wlock = threading.Lock()
        self.connected = False

    def connect(self, timeout: float = 10.0) -> None:
        deadline = time.time() + timeout
        while True:
            try:
                self.sock = socket.create_connection(self.addr, timeout=max(0.05, deadline - time.time()))
                break
            except OSError:
                if time.time() >= deadline:
                    raise ConnectTimeout(f"cannot reach {self.addr} within {timeout}s") from None
                time.sleep(0.05)
        self.sock.settimeout(None)
        self.sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        self.connected = True
        threading.Thread(target=self._reader, daemon=True).start()

    def on(self, event: str, handler: Callable) -> None:
        with self._cv:
            self.handlers[event].append(handler)
            backlog = list(self.inbox[event])
            self.inbox[event].clear()
        for m in backlog:
            handler(m)

    def _reader(self) -> None:
        while True:
            try:
                frame = _recv_frame(self.sock)
            except OSError:
                frame = None
            if frame is None:
                break
            mid, kind, obj = frame
            if kind == 1:
                with self._cv:
                    self._acks[mid] = obj
                    self._cv.notify_all()
                continue
            if not (isinstance(obj, tuple) and len(obj) == 2 and isinstance(obj[0], str)):
                break
            event, msg = obj
            with self._cv:
                hs = list(self.handlers.get(event, ()))
                if not hs:
                    self.inbox[event].append(msg)
                    self._cv.notify_all()
            for h in hs:
                h(msg)
        with self._cv:
            self.connected = False
            self._cv.notify_all()

    def wait_event(self, event: str, timeout: float = 10.0) -> Any:
        with self._cv:
            ok = self._cv.wait_for(lambda: bool(self.inbox[event]) or not self.connected, timeout)
            if not ok or not self.inbox[event]:
                raise ConnectTimeout(f"no '{event}' within {timeout}s")
            return self.inbox[event].popleft()

    def emit(self, event: str, msg: Any, ack_timeout: float = 5.0) -> Any:
        if not self.connected:
            raise AckTimeout("not connected")
        mid = next(self._ids)
        _send_frame(self.sock, self._wlock, mid, 0, (event, msg))
        with self._cv:
            ok = self._cv.wait_for(lambda: mid in self._acks or not self.connected, ack_timeout)
            if not ok or mid not in self._acks:
                raise AckTimeout(f"no ack for '{event}' within {ack_timeout}s")
            return self._acks.pop(mid)

    def close(self) -> None:
        self.connected = False
        if self.sock is not None:
            try:
                self.sock.shutdown(socket.SHUT_RDWR)
            except OSError:
                pass
            self.sock.close()
