"""Run log with the call contract of the reference's `infolog` (code/infolog.py: `init(filename, run_name, slack_url=None)`,
`log(msg, end='\\n', slack=False)`): every message goes to stdout and, once `init` has named a file, is appended there with a
millisecond timestamp.  The reference's Slack webhook is out of scope; `slack` is accepted
and ignored."""
import atexit
import sys
import time


class _RunLog(object):
	def __init__(self):
		self.sink = None

	def open(self, filename, run_name):
		self.close()
		self.sink = open(filename, 'a', encoding='utf-8')
		bar = '-' * 65
		self.sink.write('\n{0}\nStarting new {1} training run\n{0}\n'.format(bar, run_name))
		self.sink.flush()

	def write(self, msg, end):
		sys.stdout.write('{}{}'.format(msg, end))
		if self.sink is not None:
			now = time.time()
			stamp = time.strftime('%Y-%m-%d %H:%M:%S', time.localtime(now)) + '.{:03d}'.format(int(now * 1000) % 1000)
			self.sink.write('[{}]  {}\n'.format(stamp, msg))
			self.sink.flush()

	def close(self):
		if self.sink is not None:
			self.sink.close()
			self.sink = None


_log = _RunLog()
atexit.register(_log.close)


def init(filename, run_name, slack_url=None):
	_log.open(filename, run_name)


def log(msg, end='\n', slack=False):
	_log.write(msg, end)
