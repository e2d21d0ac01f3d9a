from .dummy_dataset import DummyDataset  # noqa: F401
from .items import (LABELS2ID, ID2LABELS, ChunkItem, DatasetItem, RawPreprocessor, collate_fun,  # noqa: F401
                    expand_compact_inputs)
