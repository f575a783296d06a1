"""Consumers of a decoded partition.

* `generator_df` — same class as /root/reference/cerebro_gpdb/run_da_cerebro_standalone.py:33-56
  (the Keras/tf.data consumer): iterates the result dict in insertion order and yields
  (independent_var[i], dependent_var[i]).  Works on the numpy dict of `da.input_fn` and on the CUDA
  tensor dict of `DevicePartition.as_dict()`.
* `DeviceDADataset` — the torch consumer of /root/reference/cerebro_gpdb/run_pytorchddp_da.py:40-79
  (`DADataset`) without its per-sample Python and 32 DataLoader workers: a batch of indices is
  assembled on the GPU in one launch (NHWC -> NCHW, label lookup, epoch padding by wrapping).
"""
from . import _lib
from . import segment as _segment

IMAGENET_COUNT_PER_PAR = 160160      # run_pytorchddp_da.py:32
CRITEO_COUNT_PER_PAR = 1624157       # run_pytorchddp_da.py:33


class generator_df:
    def __init__(self, df_actual_data):
        self.df_actual_data = df_actual_data
        self.length = self.get_df_length(df_actual_data)

    def get_df_length(self, df_actual_data):
        length = 0
        for _, buf in df_actual_data.items():
            assert buf['dependent_var'].shape[0] == buf['independent_var'].shape[0]
            length += buf['dependent_var'].shape[0]
        return length

    def __call__(self):
        for _, buf in self.df_actual_data.items():
            dep, ind = buf['dependent_var'], buf['independent_var']
            for i in range(dep.shape[0]):
                yield ind[i], dep[i]


class DeviceDADataset:
    """Index -> (image NCHW, label) over a `da.DevicePartition`, batched on the device.

    length: nominal epoch length (IMAGENET_COUNT_PER_PAR / CRITEO_COUNT_PER_PAR in the reference);
    in train mode indices beyond the actual sample count wrap (`i % actual_length`,
    run_pytorchddp_da.py:52-53); otherwise the length is the actual count (run_pytorchddp_da.py:78-79).
    """

    def __init__(self, partition, length=None, train=True, dataset_name='imagenet'):
        self.part = partition
        self.actual_length = len(partition)
        self.train = train
        self.dataset_name = dataset_name
        if length is None:
            length = IMAGENET_COUNT_PER_PAR if dataset_name == 'imagenet' else CRITEO_COUNT_PER_PAR
        self.length = int(length) if train else self.actual_length

    def __len__(self):
        return self.length

    def batch(self, indices, dtype=None, stream=None):
        """indices: int64 CUDA tensor [B].  Returns (x, y): x float32 (or bfloat16) [B, C, H, W] for
        image partitions / [B, F] otherwise, y int64 [B]."""
        import torch
        L = _lib.require_device()
        imgs = self.part.images
        assert indices.dtype == torch.int64 and indices.is_cuda and indices.is_contiguous()
        if not self.train:
            assert int(indices.max()) < self.actual_length
        B = indices.numel()
        if imgs.dim() == 4:
            n, h, w, c = imgs.shape
            hw, out_shape = h * w, (B, c, h, w)
        else:
            n = imgs.shape[0]
            c, hw = 1, int(imgs[0].numel())
            out_shape = (B,) + tuple(imgs.shape[1:])
        bf16 = dtype is not None and dtype == torch.bfloat16
        x = torch.empty(out_shape, dtype=torch.bfloat16 if bf16 else torch.float32, device=imgs.device)
        y = torch.empty(B, dtype=torch.int64, device=imgs.device)
        with torch.cuda.device(imgs.device):
            _lib.check(L.cds_batch_assemble(_segment._ptr(imgs), n, hw, c, _segment._ptr(self.part.labels),
                                            _segment._ptr(indices), B, _segment._ptr(x), 1 if bf16 else 0,
                                            _segment._ptr(y), _segment._stream_ptr(stream)))
        return x, y

    def __getitem__(self, i):
        import torch
        x, y = self.batch(torch.tensor([int(i)], dtype=torch.int64, device=self.part.images.device))
        return x[0], y[0]
