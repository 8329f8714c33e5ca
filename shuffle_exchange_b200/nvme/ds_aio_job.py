"""One external command of the NVMe sweep (a benchmark run, ``sync``, a page-cache flush), run without a shell.

``argv`` is executed as given — arguments are never joined into a shell string, so paths with spaces or shell
metacharacters in a sweep configuration cannot change what runs (the reference wrapper,
``deepspeed/nvme/ds_aio_job.py``, splits a joined string; the sweep's use is the same: run, capture output, fail
loudly on a non-zero exit code)."""
from __future__ import annotations

import shlex
import subprocess
from dataclasses import dataclass, field
from typing import List, Optional


@dataclass
class Job:
    cmd_line: List[str]
    output_file: Optional[str] = None      # stdout and stderr go here (None: inherit)
    work_dir: Optional[str] = None
    split_pieces: bool = True              # the sweep writes "--option value" as one piece: split those into words
    returncode: Optional[int] = field(default=None, init=False)

    # accessors kept for callers written against the reference class
    def cmd(self) -> List[str]:
        return list(self.cmd_line)

    def get_cwd(self) -> Optional[str]:
        return self.work_dir

    def run(self, verbose: bool = False) -> int:
        if self.split_pieces:
            argv = [word for piece in self.cmd_line for word in shlex.split(str(piece))]
        else:
            argv = [str(a) for a in self.cmd_line]
        if verbose:
            print("run:", argv)
        if self.output_file is None:
            done = subprocess.run(argv, cwd=self.work_dir)
        else:
            with open(self.output_file, "w") as sink:
                done = subprocess.run(argv, cwd=self.work_dir, stdout=sink, stderr=subprocess.STDOUT)
        self.returncode = done.returncode
        return done.returncode


def run_job(job: Job, verbose: bool = False) -> None:
    rc = job.run(verbose)
    if rc != 0:
        raise RuntimeError(f"{job.cmd_line!r} failed with exit code {rc}"
                           + (f" (output in {job.output_file})" if job.output_file else ""))
