"""Intra-worker schedulers (``core/harp-collective/.../schdynamic`` and ``schstatic``).

``DynamicScheduler``: N task objects = N threads sharing ONE input deque; ``submit`` / ``submit_all`` /
``start`` / ``pause`` (drain, then park) / ``pause_now`` (park as soon as possible: pause tokens go to
the *front*) / ``clean_input_queue`` / ``stop`` / ``wait_for_output`` / ``has_output`` / ``has_error``
(``schdynamic/DynamicScheduler.java:52-287``, ``TaskMonitor.java:62-98``).  Task objects keep
thread-private state.

``StaticScheduler``: one queue PER task, ``submit(task_id, input)``; tasks can submit to each other
through a ``Submitter`` (``schstatic/StaticScheduler.java:42-198``, ``TaskMonitor.java:48-218``).

On the GPU the same roles are played inside kernels (persistent CTAs pulling tiles from an atomic
counter; per-CTA accumulators merged at the end), so these host schedulers drive I/O, parsing and
the per-slice communication threads of the ``Rotator`` — exactly what the reference uses them for
outside its compute loops.
"""
from __future__ import annotations

import threading
from collections import deque
from typing import Any, Deque, Generic, List, Optional, Sequence, TypeVar

I = TypeVar("I")
O = TypeVar("O")

_PAUSE = object()
_STOP = object()


class Task(Generic[I, O]):
    """``schdynamic/Task.java:22-33`` / ``schstatic/Task.java:22-90``."""

    def run(self, inp: I) -> O:  # pragma: no cover - interface
        raise NotImplementedError

    # static-scheduler tasks get these injected
    _task_id: int = -1
    _num_tasks: int = 0
    _submitter: Optional["Submitter"] = None

    def get_task_id(self) -> int:
        return self._task_id

    def get_num_tasks(self) -> int:
        return self._num_tasks

    def get_submitter(self) -> Optional["Submitter"]:
        return self._submitter


class _Monitor(threading.Thread):
    def __init__(self, sched, task: Task, in_queue: "._Queue", out_queue: "._Queue", barrier_ev: threading.Event):
        super().__init__(daemon=True)
        self.sched, self.task, self.inq, self.outq = sched, task, in_queue, out_queue

    def run(self) -> None:
        while True:
            item = self.inq.take()
            if item is _STOP:
                return
            if item is _PAUSE:
                self.sched._on_pause()
                continue
            try:
                out = self.task.run(item)
                if out is not None:
                    self.outq.put(out)
                else:
                    self.outq.count_only()
            except Exception as e:  # noqa: BLE001
                self.sched._error = e
                self.outq.count_only()


class _Queue:
    def __init__(self):
        self.d: Deque[Any] = deque()
        self.cv = threading.Condition()
        self.produced = 0   # number of outputs (or completions) ever recorded

    def put(self, x, front: bool = False) -> None:
        with self.cv:
            (self.d.appendleft if front else self.d.append)(x)
            self.produced += 1
            self.cv.notify()

    def count_only(self) -> None:
        with self.cv:
            self.d.append(None)
            self.produced += 1
            self.cv.notify()

    def take(self):
        with self.cv:
            while not self.d:
                self.cv.wait()
            return self.d.popleft()

    def clear(self) -> int:
        with self.cv:
            n = sum(1 for x in self.d if x is not _PAUSE and x is not _STOP)
            self.d = deque(x for x in self.d if x is _PAUSE or x is _STOP)
            return n

    def __len__(self) -> int:
        with self.cv:
            return len(self.d)


class DynamicScheduler(Generic[I, O]):
    def __init__(self, tasks: Sequence[Task]):
        self.tasks: List[Task] = list(tasks)
        self.num_threads = len(self.tasks)
        self._inq = _Queue()
        self._outq = _Queue()
        self._threads: List[_Monitor] = []
        self._running = False
        self._paused = threading.Semaphore(0)
        self._resume = threading.Event()
        self._in_count = 0
        self._out_taken = 0
        self._error: Optional[BaseException] = None
        self._lock = threading.Lock()

    def get_tasks(self) -> List[Task]:
        return self.tasks

    def submit(self, inp: I) -> None:
        # inputs are only counted while running; queued-but-uncounted ones are added by start()
        # and un-counted again by pause()/pause_now() (DynamicScheduler.java:127-170)
        with self._lock:
            if self._running:
                self._in_count += 1
        self._inq.put(inp)

    def submit_all(self, inputs) -> None:
        for x in inputs:
            self.submit(x)

    def _queued_inputs(self) -> int:
        with self._inq.cv:
            return sum(1 for x in self._inq.d if x is not _PAUSE and x is not _STOP)

    def start(self) -> None:
        if self._running:
            return
        with self._lock:
            self._in_count += self._queued_inputs()
            self._running = True
        self._resume.set()
        if not self._threads:
            for t in self.tasks:
                th = _Monitor(self, t, self._inq, self._outq, self._resume)
                self._threads.append(th)
                th.start()

    def _on_pause(self) -> None:
        self._paused.release()
        self._resume.wait()

    def _park_all(self, front: bool) -> None:
        if not self._running:
            return
        self._resume.clear()
        for _ in self._threads:
            self._inq.put(_PAUSE, front=front)
        for _ in self._threads:
            self._paused.acquire()
        with self._lock:
            self._running = False
            self._in_count -= self._queued_inputs()

    def pause(self) -> None:
        """Let queued inputs drain, then park all threads."""
        self._park_all(front=False)

    def pause_now(self) -> None:
        """Park as soon as the current items finish; queued inputs stay queued."""
        self._park_all(front=True)

    def clean_input_queue(self) -> None:
        removed = self._inq.clear()
        with self._lock:
            if self._running:
                self._in_count -= removed

    def stop(self) -> None:
        if self._threads:
            if not self._running:
                self._resume.set()
            for _ in self._threads:
                self._inq.put(_STOP)
            for th in self._threads:
                th.join(timeout=30)
        self._threads = []
        self._running = False

    def has_output(self) -> bool:
        with self._lock:
            return self._out_taken < self._in_count

    def wait_for_output(self) -> Optional[O]:
        """Next output, or None once every submitted input has been accounted for."""
        while True:
            with self._lock:
                if self._out_taken >= self._in_count:
                    return None
                self._out_taken += 1
            out = self._outq.take()
            if out is not None:
                return out
            # a task returned None: counts as consumed, keep looking

    def has_error(self) -> bool:
        return self._error is not None

    submitAll, pauseNow, cleanInputQueue, waitForOutput, hasOutput, hasError, getTasks = (
        submit_all, pause_now, clean_input_queue, wait_for_output, has_output, has_error, get_tasks)


class Submitter:
    """Lets a static-scheduler task hand work to a sibling (``schstatic/Submitter.java``)."""

    def __init__(self, sched: "StaticScheduler"):
        self._sched = sched

    def submit(self, task_id: int, inp) -> None:
        self._sched.submit(task_id, inp)


class StaticScheduler(Generic[I, O]):
    def __init__(self, tasks: Sequence[Task]):
        self.tasks: List[Task] = list(tasks)
        self.num_threads = len(self.tasks)
        self._inq = [_Queue() for _ in self.tasks]
        self._outq = [_Queue() for _ in self.tasks]
        self._in_count = [0] * self.num_threads
        self._out_taken = [0] * self.num_threads
        self._threads: List[Optional[_Monitor]] = [None] * self.num_threads
        self._running = False
        self._paused = threading.Semaphore(0)
        self._resume = threading.Event()
        self._error: Optional[BaseException] = None
        self._lock = threading.Lock()
        sub = Submitter(self)
        for i, t in enumerate(self.tasks):
            t._task_id, t._num_tasks, t._submitter = i, self.num_threads, sub

    def get_task(self, task_id: int) -> Task:
        return self.tasks[task_id]

    def submit(self, task_id: int, inp: I) -> None:
        with self._lock:
            self._in_count[task_id] += 1
        self._inq[task_id].put(inp)

    def start(self) -> None:
        if self._running:
            return
        self._running = True
        self._resume.set()
        for i, t in enumerate(self.tasks):
            if self._threads[i] is None:
                th = _Monitor(self, t, self._inq[i], self._outq[i], self._resume)
                self._threads[i] = th
                th.start()

    def _on_pause(self) -> None:
        self._paused.release()
        self._resume.wait()

    def pause(self) -> None:
        if not self._running:
            return
        self._resume.clear()
        n = 0
        for i, th in enumerate(self._threads):
            if th is not None:
                self._inq[i].put(_PAUSE)
                n += 1
        for _ in range(n):
            self._paused.acquire()
        self._running = False

    def clean_input_queue(self) -> None:
        for i, q in enumerate(self._inq):
            removed = q.clear()
            with self._lock:
                self._in_count[i] -= removed

    def stop(self) -> None:
        if not self._running:
            self._resume.set()
        for i, th in enumerate(self._threads):
            if th is not None:
                self._inq[i].put(_STOP)
        for th in self._threads:
            if th is not None:
                th.join(timeout=30)
        self._threads = [None] * self.num_threads
        self._running = False

    def has_output(self, task_id: int) -> bool:
        with self._lock:
            return self._out_taken[task_id] < self._in_count[task_id]

    def wait_for_output(self, task_id: int) -> Optional[O]:
        while True:
            with self._lock:
                if self._out_taken[task_id] >= self._in_count[task_id]:
                    return None
                self._out_taken[task_id] += 1
            out = self._outq[task_id].take()
            if out is not None:
                return out

    def has_error(self, task_id: int = -1) -> bool:
        return self._error is not None

    getTask, cleanInputQueue, waitForOutput, hasOutput, hasError = (
        get_task, clean_input_queue, wait_for_output, has_output, has_error)
