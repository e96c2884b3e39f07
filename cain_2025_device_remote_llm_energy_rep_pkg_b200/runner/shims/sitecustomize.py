"""Linux fix-up for the reference's measurement loop (experiment/RunnerConfig.py:155-176).  On macOS psutil raises
ZombieProcess when the finished, un-reaped curl child is queried, which ends the loop; on Linux
`psutil.pid_exists` stays True and `memory_info()` returns zeros for a zombie, so the loop never exits
(SURVEY.md §7.1 step 0).  This patch makes `Process.memory_info` raise `psutil.ZombieProcess` (a
`NoSuchProcess`, which the reference catches at :177) for zombies — from the second query of that process on: the
first one always answers, so that a request the B200 finishes inside the loop's first 0.1 s still leaves one
sample of cpu_usage / memory_usage in the run table (an empty frame would put NaN there, RunnerConfig.py:232-234).
Loaded through PYTHONPATH by runner/launch.py; no reference file is edited."""
try:
    import psutil

    _orig = psutil.Process.memory_info

    def _memory_info(self):
        seen = getattr(self, "_b2_queried", False)
        self._b2_queried = True
        try:
            if seen and self.status() == psutil.STATUS_ZOMBIE:
                raise psutil.ZombieProcess(self.pid)
        except psutil.NoSuchProcess:
            raise
        return _orig(self)

    psutil.Process.memory_info = _memory_info
except Exception:  # psutil absent: nothing to patch
    pass
