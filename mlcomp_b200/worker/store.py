"""Dependency-free stand-ins for the rows and providers the report builders
and executors write to (the reference uses SQLAlchemy models,
mlcomp/db/models/report.py:11-88, task.py; sqlalchemy is not in this image).

Field names equal the reference's column names, so a maintainer can pass the
real ``Session`` / providers instead: the builders only call ``add``,
``bulk_save_objects``, ``update``/``commit``, ``by_name``, ``by_id`` and
``project``.
"""
import datetime
import sys


def now():
    return datetime.datetime.now(datetime.timezone.utc).replace(tzinfo=None)


class Row:
    _fields = ()

    def __init__(self, **kwargs):
        for f in self._fields:
            setattr(self, f, None)
        for k, v in kwargs.items():
            if k not in self._fields:
                raise TypeError(f'{type(self).__name__} has no column {k!r}')
            setattr(self, k, v)
        self.id = None

    def __repr__(self):
        body = ', '.join(f'{f}={getattr(self, f)!r}' for f in self._fields
                         if f != 'img' and getattr(self, f) is not None)
        return f'{type(self).__name__}({body})'


class ReportSeries(Row):          # db/models/report.py:11-24
    _fields = ('name', 'value', 'epoch', 'time', 'task', 'part', 'stage')


class ReportImg(Row):             # db/models/report.py:27-59
    _fields = ('group', 'epoch', 'task', 'img', 'dag', 'part', 'project',
               'y_pred', 'y', 'score', 'size') + \
        tuple(f'attr{i}' for i in range(1, 10)) + \
        tuple(f'attr{i}_str' for i in range(1, 6))

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.size = sys.getsizeof(self.img)      # report.py:56-59


class Report(Row):                # db/models/report.py:62-70
    _fields = ('config', 'time', 'name', 'project', 'layout')


class ReportTasks(Row):           # db/models/report.py:73-78
    _fields = ('report', 'task')


class ReportLayout(Row):          # db/models/report.py:81-86
    _fields = ('name', 'content', 'last_modified')


class Task(Row):                  # the columns this path touches, db/models/task.py
    _fields = ('name', 'status', 'executor', 'dag', 'parent', 'report',
               'score', 'result', 'additional_info', 'debug', 'gpu',
               'computer_assigned', 'loader_name', 'batch_index',
               'batch_total', 'epoch_duration', 'epoch_time_remaining',
               'current_step', 'type')


class Dag(Row):
    _fields = ('name', 'project', 'config', 'img_size')


# The two layouts the migration seeds for this path
# (mlcomp/migration/versions/002/report_layout/{img-classify,segment}.yml),
# reduced to the `items` the builders read.
LAYOUTS = {
    'img_classify': """
metric: {name: accuracy, minimize: False}
items:
  accuracy: {key: accuracy, type: series}
  img_classify: {type: img_classify, confusion_matrix: True}
""",
    'img_segment': """
metric: {name: dice, minimize: False}
items:
  dice: {key: dice, type: series}
  img_segment: {type: img_segment}
""",
}
LAYOUTS['img-classify'] = LAYOUTS['img_classify']
LAYOUTS['segment'] = LAYOUTS['img_segment']


class MemorySession:
    """In-process stand-in for mlcomp.db.core.Session + the providers."""

    def __init__(self, project=1):
        self.tables = {}
        self.project_id = project
        self.layouts = {k: ReportLayout(name=k, content=v)
                        for k, v in LAYOUTS.items()}
        self.commits = 0

    # provider surface -----------------------------------------------------
    def add(self, row):
        rows = self.tables.setdefault(type(row).__name__, [])
        row.id = len(rows) + 1
        rows.append(row)
        return row

    def bulk_save_objects(self, rows):
        for r in rows:
            self.add(r)

    def update(self):
        self.commits += 1

    commit = update

    def by_name(self, name):
        if name not in self.layouts:
            raise KeyError(f'report layout {name!r} not found')
        return self.layouts[name]

    def by_id(self, id_):
        for r in self.tables.get('Dag', []):
            if r.id == id_:
                return r
        return None

    def project(self, task_id):
        class _P:
            id = self.project_id
        return _P()

    def rows(self, cls):
        return list(self.tables.get(cls.__name__, []))

    def find_dependents(self, task_id):
        return [t for t in self.tables.get('Task', [])
                if getattr(t, '_child_of', None) == task_id]
