"""Model check (CPU, pure Python) of the tagged-record protocol used by the halo slots (multi-GPU) and by
engine 3 (lbm_device.cuh "halo slots", lbm_kernels.cuh "Engine 3").

The device code relies on one argument: a record {payload, tag} is written by ONE atomic 16-byte store;
the consumer re-reads until every record it needs carries the step's tag; a record slot is rewritten two
steps later, and "by then every reader has published the output of the step that read it, which the writer
had to see before it could compute" - so there is no flag, no fence and no acknowledgement.

The model runs that protocol under an adversarial scheduler: every record read and every record write is
its own event, events of different tasks interleave in random order, the records of one group land in
random order (a reader can see a half-written group), and readers retry exactly like the kernel (all
groups again).  It checks that every task computes exactly what a sequential reference computes and that
nobody waits forever (a slot overwritten too early would show up as a tag that can never match again).
A deliberately broken variant (single buffer instead of two parities) must be caught.
"""
import random

import pytest

GROUPS = 3          # to the ring neighbour below, to the one above, to the task itself
RECORDS = 4         # records per group (4 x 16 bytes per lane in the kernel)


def reference(n_tasks, steps, seed):
    """Sequential run: value[t] of every task after each step (a hash of its three inputs)."""
    vals = [hash((seed, t)) & 0xFFFFFFFF for t in range(n_tasks)]
    for s in range(steps):
        vals = [hash((vals[(t - 1) % n_tasks], vals[t], vals[(t + 1) % n_tasks], s)) & 0xFFFFFFFF for t in range(n_tasks)]
    return vals


class Task:
    def __init__(self, t, n, steps, value, rng, parities):
        self.t, self.n, self.steps, self.value, self.rng, self.parities = t, n, steps, value, rng, parities
        self.step = 0
        self.phase = "publish"            # prologue: publish the initial value with the first tag
        self.pending = self._writes(0, 1)
        self.got = {}
        self.done = False
        self.retries = 0

    # group g of task t is read by: 0 -> task t+1 (its "south" input), 1 -> task t-1 ("north"), 2 -> t itself
    def _writes(self, parity, tag):
        w = [(parity, self.t, g, r, (self.value, g, r), tag) for g in range(GROUPS) for r in range(RECORDS)]
        self.rng.shuffle(w)
        return w

    def _reads(self, parity):
        src = [((self.t - 1) % self.n, 0), ((self.t + 1) % self.n, 1), (self.t, 2)]
        rd = [(parity, t, g, r) for (t, g) in src for r in range(RECORDS)]
        self.rng.shuffle(rd)
        return rd

    def event(self, mem):
        """Executes ONE memory event of this task."""
        if self.phase == "publish" or self.phase == "write":
            p, t, g, r, payload, tag = self.pending.pop()
            mem[(p % self.parities, t, g, r)] = (payload, tag)
            if not self.pending:
                if self.step == self.steps:
                    self.done = True
                    return
                self.phase = "read"
                self.pending = self._reads(self.step & 1)
                self.got = {}
        elif self.phase == "read":
            key = self.pending.pop()
            self.got[key] = mem.get((key[0] % self.parities,) + key[1:], (None, 0))
            if not self.pending:
                want = 1 + self.step
                if all(tag == want for (_, tag) in self.got.values()):
                    ins = {}
                    for (p, t, g, r), (payload, _) in self.got.items():
                        assert payload[1] == g and payload[2] == r          # the record that was asked for
                        ins.setdefault(t, set()).add(payload[0])
                    assert all(len(v) == 1 for v in ins.values()), "a group mixed two steps' payloads under one tag"
                    lo, me, hi = (ins[(self.t - 1) % self.n].pop(), ins[self.t].pop(), ins[(self.t + 1) % self.n].pop()) \
                        if self.n >= 3 else (None, None, None)
                    self.value = hash((lo, me, hi, self.step)) & 0xFFFFFFFF
                    self.step += 1
                    self.phase = "write"
                    # the result goes to the OTHER parity with the next tag (the last step still publishes: the
                    # model has no SoA write-back)
                    self.pending = self._writes(self.step & 1, 1 + self.step)
                else:                      # not all there yet: poll everything again, as the kernel does
                    self.retries += 1
                    self.pending = self._reads(self.step & 1)
                    self.got = {}


def run_model(n_tasks, steps, seed, parities=2, max_events=2_000_000):
    rng = random.Random(seed)
    tasks = [Task(t, n_tasks, steps, hash((seed, t)) & 0xFFFFFFFF, rng, parities) for t in range(n_tasks)]
    mem = {}
    events = 0
    # adversarial-ish scheduler: random task, with random bursts so that some tasks run far ahead of others
    while not all(t.done for t in tasks):
        t = rng.choice(tasks)
        for _ in range(rng.choice((1, 1, 2, 5, 40))):
            if t.done:
                break
            t.event(mem)
            events += 1
        if events > max_events:
            return None, tasks             # livelock: somebody waits for a tag that can no longer appear
    return [t.value for t in tasks], tasks


@pytest.mark.parametrize("n_tasks,steps", [(3, 7), (4, 6), (5, 9), (8, 6), (16, 5)])
def test_tagged_record_protocol_is_safe_under_random_interleavings(n_tasks, steps):
    want = reference(n_tasks, steps, seed=11)
    retried = 0
    for schedule in range(40):
        got, tasks = run_model(n_tasks, steps, seed=11 + 1000 * schedule)
        assert got is not None, "a task waited forever: a slot was overwritten before its reader was done"
        # same initial values for every schedule (seeded per task), so the result must not depend on the schedule
        want = reference(n_tasks, steps, seed=11 + 1000 * schedule)
        assert got == want
        retried += sum(t.retries for t in tasks)
    assert retried > 0                     # the schedules did exercise half-written groups and late neighbours


def test_model_catches_a_broken_protocol():
    """With ONE buffer instead of two parities a fast task overwrites records its neighbour still needs:
    the model must notice (wrong result, mixed payloads, or a wait that never ends)."""
    caught = 0
    for schedule in range(40):
        try:
            got, _ = run_model(6, 6, seed=5 + 1000 * schedule, parities=1, max_events=200_000)
            if got is None or got != reference(6, 6, seed=5 + 1000 * schedule):
                caught += 1
        except AssertionError:
            caught += 1
    assert caught > 0


# ---------------------------------------------------------------------------------------------------------------
# Round 2: two more ways to break the protocol, both met on the way (profiles/r02_checked_build.md, lbm_kernels.cuh
# "record parity continues across runs").  The model below generalises the one above:
#   * split_reads: a record is read as TWO events, payload first and tag later - what ptxas made of an
#     ld.relaxed.sys.b128 whose words were only partly used.  A writer can slip in between: new tag, old payload.
#   * several runs, tasks grouped into GPUs: a GPU's tasks all finish run k before any of them starts run k+1 (kernel
#     boundary), but GPUs are not ordered against each other.  The last step of a run publishes nothing (the result goes
#     to the SoA lattice), the next run's prologue publishes with the run's first tag.  If the prologue always used
#     parity 0 ("restart"), a fast GPU's prologue of run k+1 could overwrite what a slow neighbour's LAST step of run k
#     still has to read whenever that step reads parity 0 (odd run length); with the parity continuing across runs
#     ((steps done + s) & 1) it never can.
class RunTask:
    def __init__(self, t, n, runs, value, rng, gpu, parity_rule, split_reads):
        self.t, self.n, self.runs, self.value, self.rng, self.gpu = t, n, runs, value, rng, gpu
        self.parity_rule, self.split = parity_rule, split_reads
        self.run, self.step, self.base, self.done_steps = 0, 0, 0, 0
        self.global_step = 0
        self.retries = 0
        self.finished_runs = 0
        self.done = False
        self.phase = "idle"               # waits for the scheduler's permission to start a run (kernel launch)

    def par(self, s):                     # record parity of local step s of the current run
        return ((self.done_steps + s) & 1) if self.parity_rule == "continue" else (s & 1)

    def start_run(self):
        self.step = 0
        self.phase = "write"
        self.pending = self._writes(self.par(0), self.base + 1)

    def _writes(self, parity, tag):
        w = [(parity, self.t, g, r, (self.value, g, r), tag) for g in range(GROUPS) for r in range(RECORDS)]
        self.rng.shuffle(w)
        return w

    def _reads(self, parity):
        src = [((self.t - 1) % self.n, 0), ((self.t + 1) % self.n, 1), (self.t, 2)]
        keys = [(parity, t, g, r) for (t, g) in src for r in range(RECORDS)]
        if self.split:                    # payload read and tag read are separate events, every payload before its tag
            self.rng.shuffle(keys)
            tags = [("tag", k) for k in keys]
            self.rng.shuffle(tags)
            ev = [("payload", k) for k in keys] + tags
            ev.reverse()                  # events are popped from the end: payloads first, then tags
            return ev
        rd = [("both", k) for k in keys]
        self.rng.shuffle(rd)
        return rd

    def event(self, mem):
        if self.phase == "write":
            p, t, g, r, payload, tag = self.pending.pop()
            mem[(p, t, g, r)] = (payload, tag)
            if not self.pending:
                self.phase = "read"
                self.pending = self._reads(self.par(self.step))
                self.got_payload, self.got_tag = {}, {}
        elif self.phase == "read":
            kind, key = self.pending.pop()
            payload, tag = mem.get(key, (None, 0))
            if kind in ("payload", "both"):
                self.got_payload[key] = payload
            if kind in ("tag", "both"):
                self.got_tag[key] = tag
            if not self.pending:
                want = self.base + 1 + self.step
                if all(tag == want for tag in self.got_tag.values()):
                    ins = {}
                    for (p, t, g, r), payload in self.got_payload.items():
                        assert payload is not None and payload[1] == g and payload[2] == r
                        ins.setdefault(t, set()).add(payload[0])
                    assert all(len(v) == 1 for v in ins.values()), "a group mixed two steps' payloads under one tag"
                    lo, me, hi = ins[(self.t - 1) % self.n].pop(), ins[self.t].pop(), ins[(self.t + 1) % self.n].pop()
                    self.value = hash((lo, me, hi, self.global_step)) & 0xFFFFFFFF
                    self.step += 1
                    self.global_step += 1
                    if self.step == self.runs[self.run]:       # last step of the run: result to the SoA lattice, no records
                        self.done_steps += self.runs[self.run]
                        self.base += self.runs[self.run] + 1
                        self.run += 1
                        self.finished_runs += 1
                        self.phase = "idle"
                        self.done = self.run == len(self.runs)
                    else:
                        self.phase = "write"
                        self.pending = self._writes(self.par(self.step), self.base + 1 + self.step)
                else:
                    self.retries += 1
                    self.pending = self._reads(self.par(self.step))
                    self.got_payload, self.got_tag = {}, {}


def run_model_runs(n_tasks, runs, seed, gpus=2, parity_rule="continue", split_reads=False, max_events=400_000):
    assert n_tasks >= 3
    rng = random.Random(seed)
    tasks = [RunTask(t, n_tasks, runs, hash((seed, t)) & 0xFFFFFFFF, rng, t * gpus // n_tasks, parity_rule, split_reads)
             for t in range(n_tasks)]
    mem, events = {}, 0
    while not all(t.done for t in tasks):
        t = rng.choice(tasks)
        if t.phase == "idle" and not t.done:
            # kernel boundary: a GPU launches run k only when all ITS tasks have finished run k-1
            if all(o.finished_runs >= t.run for o in tasks if o.gpu == t.gpu):
                t.start_run()
            else:
                continue
        for _ in range(rng.choice((1, 1, 2, 5, 40))):
            if t.done or t.phase == "idle":
                break
            t.event(mem)
            events += 1
        if events > max_events:
            return None, tasks
    return [t.value for t in tasks], tasks


def reference_runs(n_tasks, total_steps, seed):
    vals = [hash((seed, t)) & 0xFFFFFFFF for t in range(n_tasks)]
    for s in range(total_steps):
        vals = [hash((vals[(t - 1) % n_tasks], vals[t], vals[(t + 1) % n_tasks], s)) & 0xFFFFFFFF for t in range(n_tasks)]
    return vals


@pytest.mark.parametrize("runs", [[3, 3], [1, 4, 1, 3], [5, 2, 5], [1, 1, 1, 1]])
@pytest.mark.parametrize("n_tasks,gpus", [(4, 2), (6, 3), (8, 2), (3, 3)])
def test_several_runs_across_gpus_are_safe_when_the_record_parity_continues(n_tasks, gpus, runs):
    for schedule in range(25):
        seed = 7 + 1000 * schedule
        got, _ = run_model_runs(n_tasks, runs, seed, gpus=gpus, parity_rule="continue")
        assert got is not None, "a task waited forever"
        assert got == reference_runs(n_tasks, sum(runs), seed)


def _caught(**kw):
    """In how many of 60 random schedules (6 tasks on 3 GPUs, three runs of 3 steps) the model notices a fault."""
    caught = 0
    for schedule in range(60):
        seed = 3 + 1000 * schedule
        try:
            got, _ = run_model_runs(6, [3, 3, 3], seed, gpus=3, max_events=150_000, **kw)
            if got is None or got != reference_runs(6, 9, seed):
                caught += 1
        except AssertionError:
            caught += 1
    return caught


def test_model_catches_a_prologue_that_restarts_the_record_parity():
    """Odd run lengths: the last step of a run reads parity 0, which a faster GPU's next prologue would overwrite
    (59 of 60 schedules show it; the shipped rule none)."""
    assert _caught(parity_rule="continue") == 0
    assert _caught(parity_rule="restart") > 30


def test_model_catches_record_loads_split_into_payload_and_tag():
    """What ptxas did to ld.relaxed.sys.b128 in the self-checking build: payload word first, tag word later."""
    assert _caught(parity_rule="continue", split_reads=True) > 30
