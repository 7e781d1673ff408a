"""Binds the delta wire format (core/wire.py) to the reference's transport.

The reference ships one step to the workers as ``Executor.submit("execute_model", batch)``: the whole
``ForwardBatch`` is pickled into every worker's ``mp.Queue`` (core/executor.py:86-100,
core/worker_client.py:51-52) and unpickled in ``WorkerClient.worker_main_loop`` (:87-94).  ``install()`` keeps
that transport (same queues, same request envelope, same method name) and changes only the payload:

* engine process: ``Executor.execute_model`` encodes the batch ONCE with a ``BatchEncoder`` and submits the
  bytes (the same bytes go to every tensor-parallel rank, as the reference sends the same batch to all workers);
  ``Scheduler.finish_sequence`` / ``Scheduler.preempt`` tell the encoder to forget the sequence
  (core/scheduler.py:140-153, 173-190), so a re-admitted sequence is resent in full;
* worker process: ``Worker.execute_model`` accepts bytes, applies them to its ``BatchDecoder`` (mirror
  ``Sequence`` objects of the reference's own class) and runs the unchanged model runner on the rebuilt
  ``ForwardBatch``.  A lost or corrupted message raises in the worker instead of desynchronising it.  Handles the
  scheduler drops (finish / preempt) and sequences it resends in full also make the worker's attention backend
  forget its device slot row (``SlotTable.forget``): the finish signal the pickle transport does not have.
"""
from __future__ import annotations

_INSTALLED = False


def install(verify: bool = False) -> None:
    """Idempotent; must run in the engine process AND in every worker process (the harness applies it at import
    time in each spawned process)."""
    global _INSTALLED
    if _INSTALLED:
        return
    import nanovllmx.core.executor as E
    import nanovllmx.core.scheduler as S
    import nanovllmx.core.worker as W
    from nanovllmx.core.common import ForwardBatch, ForwardMode, Sequence

    from .wire import BatchDecoder, BatchEncoder

    encoders: list = []          # the engine process has one Executor; kept in a list so the scheduler hooks find it

    def execute_model(self, batch):
        enc = getattr(self, "_b200_encoder", None)
        if enc is None:
            enc = self._b200_encoder = BatchEncoder(verify=verify)
            encoders.append(enc)
        return self.submit("execute_model", enc.encode(batch))

    orig_finish, orig_preempt = S.Scheduler.finish_sequence, S.Scheduler.preempt

    def finish_sequence(self, seq):
        for enc in encoders:
            enc.forget(seq.seq_id)
        return orig_finish(self, seq)

    def preempt(self, seq):
        for enc in encoders:
            enc.forget(seq.seq_id)
        return orig_preempt(self, seq)

    orig_worker_execute = W.Worker.execute_model

    def worker_execute_model(self, batch):
        if isinstance(batch, (bytes, bytearray, memoryview)):
            dec = getattr(self, "_b200_decoder", None)
            if dec is None:
                def on_drop(seq_id, worker=self):
                    # finish / preempt / full resend: the attention backend's device row of that id is stale from now
                    # on (the stock FlashInfer backend keeps no rows and has no slot_table)
                    table = getattr(getattr(getattr(worker, "model_runner", None), "attn_backend", None), "slot_table", None)
                    if table is not None:
                        table.forget(seq_id)
                dec = self._b200_decoder = BatchDecoder(sequence_cls=Sequence, batch_cls=ForwardBatch, mode_cls=ForwardMode,
                                                        on_drop=on_drop)
            batch = dec.decode(batch)
        return orig_worker_execute(self, batch)

    E.Executor.execute_model = execute_model
    S.Scheduler.finish_sequence = finish_sequence
    S.Scheduler.preempt = preempt
    W.Worker.execute_model = worker_execute_model
    _INSTALLED = True
