"""contrib/transform/tta.py:9-31 without albumentations: transforms are the
short tags of utils/config.py:93-104 ('hflip', 'vflip', 'transpose');
albumentations objects are accepted too and mapped by class name."""
from mlcomp_b200 import ops

_BY_CLASS = {'HorizontalFlip': 'hflip', 'VerticalFlip': 'vflip',
             'Transpose': 'transpose'}


def parse_albu_short(config, always_apply=False):
    """utils/config.py:93-104 reduced to the tag vocabulary: the short string
    form returns the tag itself; the dict form (``{'name':
    'HorizontalFlip', ...}``, which the reference hands to ``parse_albu``,
    :77-90) is accepted for the three transforms TtaWrap can invert
    (tta.py:22-29) and for a ``Compose`` of them (-> list of tags).  Any other
    albumentations class cannot be built here (albumentations is not a
    dependency of this package) and raises."""
    if isinstance(config, str):
        if config in ('hflip', 'vflip', 'transpose'):
            return config
        raise Exception(f'Unknwon augmentation {config}')
    assert type(config) == dict
    assert 'name' in config, f'name is required in {config}'
    name = config['name']
    if name == 'Compose':
        return [parse_albu_short(c, always_apply)
                for c in config.get('items', [])]
    if name in _BY_CLASS:
        return _BY_CLASS[name]
    raise Exception(f'augmentation {name} has no test-time inverse '
                    '(TtaWrap undoes HorizontalFlip, VerticalFlip, Transpose)')


class TtaWrap:
    def __init__(self, dataset, tfms=()):
        self.dataset = dataset
        self.tfms = tfms
        self.tags = []
        for t in tfms:
            tag = t if isinstance(t, str) else _BY_CLASS.get(type(t).__name__)
            if tag is not None:      # tta.py:22-29 ignores anything else
                self.tags.append(parse_albu_short(tag))

    def __getitem__(self, item):
        return self.dataset[item]

    def __len__(self):
        return len(self.dataset)

    def inverse(self, a):
        """Materialised inverse (tta.py:20-31) for callers that want the
        tensor; the scoring kernels never need it (they fold ``self.tags``
        into their load addresses)."""
        if a.dim() not in (3, 4):
            raise ValueError('inverse expects a 3-D or 4-D tensor')
        x = a.contiguous()
        x = x[None] if a.dim() == 3 else x
        out = ops.seg_prob([x], [self.tags], None).to(a.dtype)
        return out[0] if a.dim() == 3 else out


__all__ = ['TtaWrap', 'parse_albu_short']
