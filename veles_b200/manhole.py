"""Manhole: an interactive Python shell inside a running process, reachable through a UNIX socket
(``/tmp/manhole-<pid>``).  ``--manhole`` installs it at start, SIGUSR2 installs it on demand
(reference: ``veles/thread_pool.py:139-164, 527-533``, which delegates to the third-party ``manhole`` package).

    $ python -m veles_b200.manhole <pid>        # or: socat - UNIX-CONNECT:/tmp/manhole-<pid>
    >>> workflow.decision.epoch_number
    >>> launcher.pause(); launcher.resume()

One thread accepts connections, each connection gets its own ``code.InteractiveConsole`` whose namespace holds
``workflow``, ``launcher`` and ``dump_stacks()``.  The socket is owned by the user (mode 0600): whoever can open it can
run code as this process, exactly like the reference's manhole.
"""
from __future__ import annotations

import code
import io
import os
import socket
import sys
import threading
import traceback

_installed = {}
_lock = threading.Lock()


def socket_path(pid=None):
    return "/tmp/manhole-%d" % (pid or os.getpid())


def dump_stacks(file=None):
    """Print the stack of every thread (what SIGUSR1 prints)."""
    out = file or sys.stderr
    names = {t.ident: t.name for t in threading.enumerate()}
    for ident, frame in sys._current_frames().items():
        print("Thread %s (%s):" % (ident, names.get(ident, "?")), file=out)
        traceback.print_stack(frame, file=out)
    return None


class _Console(code.InteractiveConsole):
    def __init__(self, conn, namespace):
        code.InteractiveConsole.__init__(self, namespace)
        self.conn = conn
        self.rfile = conn.makefile("r", encoding="utf-8", errors="replace")

    def write(self, data):
        try:
            self.conn.sendall(data.encode("utf-8", errors="replace"))
        except OSError:
            pass

    def raw_input(self, prompt=""):
        self.write(prompt)
        line = self.rfile.readline()
        if not line:
            raise EOFError()
        return line.rstrip("\n")

    def runcode(self, codeobj):
        # print() / displayhook output of the executed statement goes to the client
        old_out, old_err = sys.stdout, sys.stderr
        buf = io.StringIO()
        sys.stdout = sys.stderr = buf
        try:
            code.InteractiveConsole.runcode(self, codeobj)
        finally:
            sys.stdout, sys.stderr = old_out, old_err
        text = buf.getvalue()
        if text:
            self.write(text)


def _serve(server, namespace):
    while True:
        try:
            conn, _ = server.accept()
        except OSError:
            return
        threading.Thread(target=_session, args=(conn, namespace), name="manhole-session", daemon=True).start()


def _session(conn, namespace):
    console = _Console(conn, dict(namespace))
    try:
        console.interact(banner="veles_b200 manhole (pid %d); workflow, launcher, dump_stacks() are defined"
                         % os.getpid(), exitmsg="")
    except (OSError, EOFError, SystemExit):
        pass
    finally:
        try:
            conn.close()
        except OSError:
            pass


def install(workflow=None, launcher=None, path=None, **extra):
    """Start the manhole (idempotent per path); returns the socket path."""
    path = path or socket_path()
    with _lock:
        if path in _installed:
            _installed[path]["namespace"].update(workflow=workflow, launcher=launcher, **extra)
            return path
        try:
            os.unlink(path)
        except OSError:
            pass
        server = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        old = os.umask(0o177)
        try:
            server.bind(path)
        finally:
            os.umask(old)
        server.listen(2)
        namespace = dict(workflow=workflow, launcher=launcher, dump_stacks=dump_stacks, **extra)
        t = threading.Thread(target=_serve, args=(server, namespace), name="manhole", daemon=True)
        _installed[path] = {"server": server, "thread": t, "namespace": namespace}
        t.start()
    return path


def uninstall(path=None):
    path = path or socket_path()
    with _lock:
        entry = _installed.pop(path, None)
    if entry is not None:
        try:
            entry["server"].close()
        except OSError:
            pass
        try:
            os.unlink(path)
        except OSError:
            pass


def connect(pid_or_path, commands=None, timeout=5.0):
    """Client: run ``commands`` (a list of source lines) and return the transcript, or attach stdin/stdout."""
    path = pid_or_path if isinstance(pid_or_path, str) and not pid_or_path.isdigit() else socket_path(int(pid_or_path))
    s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    s.settimeout(timeout)
    s.connect(path)
    if commands is not None:
        for line in commands:
            s.sendall((line + "\n").encode())
        s.shutdown(socket.SHUT_WR)
        chunks = []
        try:
            while True:
                d = s.recv(65536)
                if not d:
                    break
                chunks.append(d)
        except socket.timeout:
            pass
        s.close()
        return b"".join(chunks).decode(errors="replace")
    s.settimeout(None)

    def pump():
        while True:
            d = s.recv(65536)
            if not d:
                os._exit(0)
            sys.stdout.write(d.decode(errors="replace"))
            sys.stdout.flush()
    threading.Thread(target=pump, daemon=True).start()
    for line in sys.stdin:
        s.sendall(line.encode())
    s.close()
    return None


def main(argv=None):
    """``veles_manhole <pid | socket path>``: attach a terminal to the REPL of a running workflow."""
    argv = sys.argv[1:] if argv is None else list(argv)
    if len(argv) != 1 or argv[0] in ("-h", "--help"):
        print("usage: python -m veles_b200.manhole <pid | socket path>")
        return 0 if argv else 2
    connect(argv[0])
    return 0


if __name__ == "__main__":
    sys.exit(main())
