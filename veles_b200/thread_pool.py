"""Worker thread pool for the dataflow scheduler.

Parity target: ``veles/thread_pool.py:71-608`` (a Twisted ``ThreadPool`` subclass with
pause/resume, thread enter/exit and shutdown callback sets, global SIGINT handling, SIGUSR1
stack dumps, ``errback`` that stops everything on a worker exception).  Twisted is not used
here: this is a small queue-based pool on ``threading`` with the same observable behaviour.
On the B200 path GPU work is asynchronous, so two host threads (the reference default,
``veles/config.py:253-256``) are plenty to keep the launch queue full.
"""
from __future__ import annotations

import queue
import signal
import sys
import threading
import traceback
import weakref

from .logger import Logger


class ThreadPool(Logger):
    pools = weakref.WeakSet()
    _sigint_installed = False
    _old_sigint = None

    def __init__(self, minthreads=2, maxthreads=2, name=None, **kwargs):
        super(ThreadPool, self).__init__(**kwargs)
        self.name = name or "pool"
        self.minthreads = max(1, minthreads)
        self.maxthreads = max(self.minthreads, maxthreads)
        self.q = queue.Queue()
        self.workers = []
        self.started = False
        self.joined = False
        self.failure = None
        self.silent = False
        self._not_paused = threading.Event()
        self._not_paused.set()
        self._lock = threading.Lock()
        self._busy = 0
        self._idle_cond = threading.Condition(self._lock)
        self.on_shutdowns = []
        self.on_thread_enters = []
        self.on_thread_exits = []
        self.on_failures = []
        ThreadPool.pools.add(self)

    # -- lifecycle ----------------------------------------------------------------------
    def start(self):
        with self._lock:
            if self.started:
                return
            self.started = True
            self.joined = False
            for i in range(self.minthreads):
                self._spawn(i)

    def _spawn(self, i):
        t = threading.Thread(target=self._worker, name="%s-%d" % (self.name, i), daemon=True)
        self.workers.append(t)
        t.start()

    def _worker(self):
        for cb in list(self.on_thread_enters):
            self._safe(cb)
        while True:
            item = self.q.get()
            if item is None:
                break
            fn, args, kwargs = item
            self._not_paused.wait()
            try:
                if self.failure is None:
                    fn(*args, **kwargs)
            except BaseException as e:  # noqa: BLE001 - reported via errback
                self.errback(e)
            finally:
                with self._lock:
                    self._busy -= 1
                    if self._busy == 0:
                        self._idle_cond.notify_all()
        for cb in list(self.on_thread_exits):
            self._safe(cb)

    def _safe(self, cb):
        try:
            cb()
        except Exception:  # noqa: BLE001
            self.exception("thread pool callback failed")

    def callInThread(self, fn, *args, **kwargs):
        if self.joined:
            return
        if not self.started:
            self.start()
        with self._lock:
            self._busy += 1
            if self._busy > len(self.workers) and len(self.workers) < self.maxthreads:
                self._spawn(len(self.workers))
        self.q.put((fn, args, kwargs))

    call_in_thread = callInThread

    def wait_idle(self, timeout=None):
        with self._lock:
            return self._idle_cond.wait_for(lambda: self._busy == 0, timeout)

    def pause(self):
        self._not_paused.clear()

    def resume(self):
        self._not_paused.set()

    @property
    def paused(self):
        return not self._not_paused.is_set()

    def errback(self, failure):
        """Any exception in a worker stops the pool and notifies listeners."""
        if self.failure is None:
            self.failure = failure
        if not self.silent:
            self.error("Unhandled exception in a worker thread:\n%s",
                       "".join(traceback.format_exception(type(failure), failure, failure.__traceback__)))
        for cb in list(self.on_failures):
            try:
                cb(failure)
            except Exception:  # noqa: BLE001
                pass

    def register_on_shutdown(self, func, weak=True):
        self.on_shutdowns.append(weakref.WeakMethod(func) if weak and hasattr(func, "__self__") else (lambda f=func: f))

    def unregister_on_shutdown(self, func):
        self.on_shutdowns = [r for r in self.on_shutdowns if r() is not func and r() != func]

    def shutdown(self, execute_remaining=False, force=False, timeout=0.25):
        with self._lock:
            if self.joined:
                return
            self.joined = True
        for ref in list(self.on_shutdowns):
            cb = ref()
            if cb is not None:
                self._safe(cb)
        self.on_shutdowns = []
        if not execute_remaining:
            try:
                while True:
                    item = self.q.get_nowait()
                    if item is not None:
                        with self._lock:
                            self._busy -= 1
            except queue.Empty:
                pass
        self._not_paused.set()
        for _ in self.workers:
            self.q.put(None)
        me = threading.current_thread()
        for t in self.workers:
            if t is not me:
                t.join(None if execute_remaining and not force else timeout)
        self.workers = []
        self.started = False

    stop = shutdown

    # -- process-wide helpers ---------------------------------------------------------------
    @staticmethod
    def shutdown_pools(execute_remaining=False):
        for pool in list(ThreadPool.pools):
            pool.shutdown(execute_remaining=execute_remaining)

    @staticmethod
    def sigint_handler(signum, frame):
        ThreadPool.shutdown_pools()
        old = ThreadPool._old_sigint
        if callable(old):
            old(signum, frame)
        else:
            raise KeyboardInterrupt()

    @staticmethod
    def install_signal_handlers():
        if threading.current_thread() is not threading.main_thread():
            return
        if not ThreadPool._sigint_installed:
            ThreadPool._old_sigint = signal.signal(signal.SIGINT, ThreadPool.sigint_handler)
            ThreadPool._sigint_installed = True
        if hasattr(signal, "SIGUSR1"):
            signal.signal(signal.SIGUSR1, ThreadPool.sigusr1_handler)
        if hasattr(signal, "SIGUSR2"):
            signal.signal(signal.SIGUSR2, ThreadPool.sigusr2_handler)

    manhole_context = {}    # what the manhole's namespace shows: set by the launcher (workflow=..., launcher=...)

    @staticmethod
    def sigusr2_handler(signum, frame):
        """Open the manhole on demand (reference: veles/thread_pool.py:527-533)."""
        from . import manhole
        path = manhole.install(**ThreadPool.manhole_context)
        print("SIGUSR2: manhole listening on %s" % path, file=sys.stderr)

    @staticmethod
    def sigusr1_handler(signum, frame):
        """Dump the stacks of all threads (reference: veles/thread_pool.py:520-546)."""
        print(ThreadPool.dump_stacks(), file=sys.stderr)

    @staticmethod
    def dump_stacks():
        names = {t.ident: t.name for t in threading.enumerate()}
        out = []
        for tid, stack in sys._current_frames().items():
            out.append("Thread %s (%s):" % (names.get(tid, "?"), tid))
            out.append("".join(traceback.format_stack(stack)))
        return "\n".join(out)

    @staticmethod
    def debug_deadlocks():
        return ThreadPool.dump_stacks()

    def __getstate__(self):
        return {"name": self.name, "minthreads": self.minthreads, "maxthreads": self.maxthreads}

    def __setstate__(self, state):
        self.__init__(**state)


def errback(failure, pool=None):
    if pool is not None:
        pool.errback(failure)
