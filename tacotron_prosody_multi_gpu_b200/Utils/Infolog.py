"""``log`` of the reference (Utils/Infolog.py:27-37) without the Slack hook (out of scope)."""
import sys
import time

_file = None


def init(filename, run_name=None, slack_url=None):
    global _file
    _file = open(filename, "a", encoding="utf-8")
    _file.write("\n-----------------------------------------------------------------\n")
    _file.write("Starting new {} training run\n".format(run_name))


def log(msg, end="\n", slack=False):
    print(msg, end=end)
    sys.stdout.flush()
    if _file is not None:
        _file.write("[%s]  %s\n" % (time.strftime("%Y-%m-%d %H:%M:%S"), msg))
        _file.flush()
