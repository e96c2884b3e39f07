"""Linux fix-up for the reference's measurement loop (experiment/RunnerConfig.py:155-176).  On macOS psutil raises
ZombieProcess when the finished, un-reaped curl child is queried, which ends the loop; on Linux
`psutil.pid_exists` stays True and `memory_info()` returns zeros for a zombie, so the loop never exits
(SURVEY.md §7.1 step 0).  This patch makes `Process.memory_info` raise `psutil.ZombieProcess` (a
`NoSuchProcess`, which the reference catches at :177) for zombies.  Loaded through PYTHONPATH by runner/launch.py;
no reference file is edited."""
try:
    import psutil

    _orig = psutil.Process.memory_info

    def _memory_info(self):
        try:
            if self.status() == psutil.STATUS_ZOMBIE:
                raise psutil.ZombieProcess(self.pid)
        except psutil.NoSuchProcess:
            raise
        return _orig(self)

    psutil.Process.memory_info = _memory_info
except Exception:  # psutil absent: nothing to patch
    pass
