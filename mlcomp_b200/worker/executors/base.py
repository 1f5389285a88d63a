"""Executor plugin API, as mlcomp/worker/executors/base/executor.py:69-227 and
step.py define it: class registry addressed from YAML ``type:`` by
``ClassName`` / ``classname`` / ``class_name``, construction from the executor
mapping, ``__call__(task=, task_provider=, dag=)`` -> ``work()``, progress and
logging helpers.  The DB-backed parts (steps, data sync) are reduced to what
runs without SQLAlchemy; the surface a subclass sees is unchanged.
"""
import re
import time
from abc import ABC, abstractmethod

import yaml
from tqdm import tqdm

_first_cap = re.compile('(.)([A-Z][a-z]+)')
_all_cap = re.compile('([a-z0-9])([A-Z])')


def to_snake(name):                      # mlcomp/utils/misc.py:145-147
    return _all_cap.sub(r'\1_\2', _first_cap.sub(r'\1_\2', name)).lower()


class StepWrap:
    """Hierarchical steps (worker/executors/base/step.py:8-135) kept in
    memory: start(level, name, index) / end(level) and levelled logging."""

    def __init__(self, session, logger, logger_db, task, task_provider):
        self.task = task
        self.task_provider = task_provider
        self.logger = logger
        self.children = []
        self.log = []

    def enter(self):
        if not self.children:
            self.start(0, 'main', 0)

    def start(self, level, name=None, index=None):
        while self.children and self.children[-1]['level'] >= level:
            self.children.pop()
        step = {'level': level, 'name': name, 'index': index or 0,
                'started': time.time()}
        self.children.append(step)
        if self.task is not None:
            self.task.current_step = '.'.join(
                str(c['index'] + 1) for c in self.children[1:])
        return step

    def end(self, level):
        while self.children and self.children[-1]['level'] >= level:
            self.children.pop()

    def _emit(self, level, message, db):
        self.log.append((level, message))
        if self.logger is not None:
            getattr(self.logger, level)(message)

    def debug(self, message, db=False):
        self._emit('debug', message, db)

    def info(self, message, db=False):
        self._emit('info', message, db)

    def warning(self, message, db=False):
        self._emit('warning', message, db)

    def error(self, message, db=False):
        self._emit('error', message, db)


class TqdmWrapper:                       # executor.py:16-66
    def __init__(self, executor, iterable=None, desc='progress', interval=10,
                 **kwargs):
        self.desc = desc
        self.interval = interval
        self.executor = executor
        self.tqdm = tqdm(iterable=iterable, **kwargs)

    def refresh(self):
        task, bar = self.executor.task, self.tqdm
        if task is not None:
            task.loader_name = self.desc
            task.batch_index = bar.n
            task.batch_total = bar.total
            task.epoch_duration = time.time() - bar.start_t
            if bar.n > 0 and bar.total:
                task.epoch_time_remaining = task.epoch_duration * \
                    (bar.total - bar.n) / bar.n
            if self.executor.task_provider is not None:
                self.executor.task_provider.update()
        return time.time()

    def set_description(self, desc=None, refresh=True):
        self.desc = desc or ''
        if refresh:
            self.refresh()

    def __iter__(self):
        last = self.refresh()
        for item in self.tqdm:
            if time.time() - last > self.interval:
                last = self.refresh()
            yield item
        self.refresh()


class Executor(ABC):
    _child = dict()

    session = None
    task_provider = None
    logger = None
    logger_db = None
    step = None
    task = None
    dag = None

    def __init__(self, **kwargs):
        self.kwargs = kwargs

    # logging (executor.py:81-103) ------------------------------------------
    def _log(self, level, message, db):
        if self.step:
            getattr(self.step, level)(message, db=db)
        else:
            print(message)

    def debug(self, message, db=False):
        self._log('debug', message, db)

    def info(self, message, db=False):
        self._log('info', message, db)

    def warning(self, message, db=False):
        self._log('warning', message, db)

    def error(self, message, db=False):
        self._log('error', message, db)

    def write(self, message):
        if message.strip() != '':
            self.info(message, db=True)

    def flush(self):
        pass

    # invocation (executor.py:119-139) ---------------------------------------
    def __call__(self, *, task, task_provider, dag):
        assert dag is not None, 'You must fetch task with dag_rel'
        self.task_provider = task_provider
        self.task = task
        self.dag = dag
        self.step = StepWrap(self.session, self.logger, self.logger_db, task,
                             task_provider)
        self.step.enter()
        res = self.work()
        if task_provider is not None:
            task_provider.commit()
        return res

    @abstractmethod
    def work(self) -> dict:
        pass

    # construction (executor.py:141-177) -------------------------------------
    @classmethod
    def _from_config(cls, executor: dict, config, additional_info: dict):
        return cls(**executor)

    @staticmethod
    def from_config(*, executor: str, config, additional_info: dict,
                    session=None, logger=None, logger_db=None) -> 'Executor':
        if executor not in config['executors']:
            raise ModuleNotFoundError(
                f'Executor {executor} has not been found')
        spec = additional_info['executor'] if additional_info and \
            'executor' in additional_info else config['executors'][executor]
        child = Executor._child[spec['type']]
        res = child._from_config(spec, config, additional_info)
        res.session = session
        res.logger = logger
        res.logger_db = logger_db
        return res

    @staticmethod
    def register(cls):
        Executor._child[cls.__name__] = cls
        Executor._child[cls.__name__.lower()] = cls
        Executor._child[to_snake(cls.__name__)] = cls
        return cls

    @staticmethod
    def is_registered(cls: str):
        return cls in Executor._child

    def tqdm(self, iterable=None, desc='progress', interval=10, **kwargs):
        return TqdmWrapper(executor=self, iterable=iterable, desc=desc,
                           interval=interval, **kwargs)

    def dependent_results(self):         # executor.py:222-227
        res = dict()
        for t in self.task_provider.find_dependents(self.task.id):
            res[t.id] = yaml.safe_load(t.result) if t.result else None
        return res


__all__ = ['Executor', 'StepWrap', 'TqdmWrapper', 'to_snake']
