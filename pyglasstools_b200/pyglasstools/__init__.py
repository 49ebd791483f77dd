"""pyglasstools -- drop-in Python API of PyGlassTools' Irving-Kirkwood / thermo path on B200.

Same module-level surface as the reference package (pyglasstools/__init__.py:16-138): ``comm, rank, size,
loggers_list, solvers_list, set_sysdata/get_sysdata/update_sysdata, set_potential/get_potential, reset,
analyze`` -- so ``example_scripts/{thermo,irvingkirkwood}`` run unchanged.  Every pair sum is executed by
``libpgt_cuda.so`` (hand-written sm_100a kernels behind the C-ABI of include/pgt_cuda.h); there is no CPU
path.  One process drives one GPU; under ``torchrun`` the ranks take the reference's MPI ranks' place.
"""
import atexit
import signal
import sys

from pyglasstools import _pyglasstools

# one communicator per process, created at import like the reference does (__init__.py:16-18)
comm = _pyglasstools.Communicator()
rank = comm.getRank()
size = comm.getSizeGlobal()

savemode = "unordered"


def set_savemode(name):
    # the reference assigns a local here (__init__.py:22-23), so the mode never changes; kept as is
    savemode = name  # noqa: F841


_previous_excepthook = sys.excepthook


def _abort_on_uncaught(exc_type, value, traceback):
    """Uncaught exception on any rank ends the whole job (__init__.py:29-35)."""
    _previous_excepthook(exc_type, value, traceback)
    sys.stdout.flush()
    sys.stderr.flush()
    comm.abort(1)


sys.excepthook = _abort_on_uncaught

loggers_list = []
solvers_list = []
globalsysdata = None
globalpotential = None


def set_sysdata(sysdata):
    global globalsysdata
    globalsysdata = sysdata


def update_sysdata(frame_num):
    globalsysdata.update(frame_num)


def get_sysdata():
    return globalsysdata


def set_potential(potential):
    global globalpotential
    globalpotential = potential


def get_potential():
    return globalpotential


def reset():
    """Forget every registered solver and logger (also run at interpreter exit)."""
    loggers_list.clear()
    solvers_list.clear()


atexit.register(reset)


class DelayedInterrupt(object):
    """Defers SIGINT / SIGTERM while the loggers write a frame, so a log never ends mid-frame
    (__init__.py:88-102)."""

    _signals = (signal.SIGINT, signal.SIGTERM)

    def __enter__(self):
        self.pending = None
        self.saved = {}
        for sig in self._signals:
            try:
                self.saved[sig] = signal.signal(sig, self._remember)
            except ValueError:          # not the main thread: nothing to defer
                pass
        return self

    def _remember(self, sig, frame):
        self.pending = (sig, frame)
        print("Signal {} received by Process [{}]. Delaying KeyboardInterrupt.".format(sig, rank))

    def __exit__(self, exc_type, value, traceback):
        for sig, handler in self.saved.items():
            signal.signal(sig, handler)
        if self.pending is not None:
            handler = self.saved.get(self.pending[0])
            if callable(handler):
                handler(*self.pending)
        return False


from pyglasstools import utils  # noqa: E402,F401


def _progressbar(total, initial):
    try:
        from tqdm import tqdm
        return tqdm(total=total, file=sys.stdout, leave=False, initial=initial)
    except ImportError:      # tqdm is cosmetic
        return None


def analyze(frame_list):
    """Frame loop of the reference (__init__.py:107-138): for each frame every solver is updated and run,
    every logger saves, rank 0 rewrites the checkpoint file with the next frame number."""
    sysdata = globalsysdata
    total = len(frame_list)
    frame_list = sysdata.setup_checkpoint(frame_list)
    if not frame_list:
        if rank == 0:
            print("No more frames to analyze. Check your logfiles/output")
        return
    bar = _progressbar(total, frame_list[0]) if rank == 0 else None
    if rank == 0:
        print("")
    for frame_num in frame_list:
        for solver in solvers_list:
            solver.update(frame_num)
            solver.run()
        with DelayedInterrupt():
            for logger in loggers_list:
                logger.save(frame_num)
            if rank == 0 and sysdata.checkpointfile is not None:
                with open(sysdata.checkpointfile, "w") as f:
                    f.write("Frame {}".format(frame_num + 1))
            if bar is not None:
                bar.update(1)
                print("")
            comm.barrier()
    if bar is not None:
        bar.close()
