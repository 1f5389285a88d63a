from .base import Executor, StepWrap, TqdmWrapper, to_snake
from .equation import Equation
from .valid import Valid
from .infer import Infer
from .segment import ValidSegment, InferSegment
from .classify import ValidClassify, InferClassify

__all__ = ['Executor', 'StepWrap', 'TqdmWrapper', 'to_snake', 'Equation',
           'Valid', 'Infer', 'ValidSegment', 'InferSegment', 'ValidClassify',
           'InferClassify']
