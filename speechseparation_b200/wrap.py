"""Signal wrap around the model, on the GPU through the C ABI.

Host-side mirror of what every caller of the reference repeats around ``BSRNN.forward``:

  * ``stft_pack``    <- m_dataset.py:187-190 / infer.py:29-33
  * ``unpack_istft`` <- m_dataset.py:192-195 / infer.py:35-37
  * ``infer``        <- m_dataset.py:182-199  (the forward-only part: x, x_time)
  * ``StreamingSeparator`` <- the per-chunk loop of infer-streaming.py:84-147
                              (= SpeechSeparator::new_chunk, speech-ladspa-onnx.cpp:171-267)
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib
from .bsrnn import BSRNN, _p, _stream_ptr

N_FFT = 2048
HOP = 1024


def _check_wav(wav: torch.Tensor):
    if not isinstance(wav, torch.Tensor) or not wav.is_cuda:
        raise RuntimeError("wav must be a CUDA tensor: speechseparation_b200 has no CPU path")
    if wav.dtype != torch.float32 or wav.dim() != 2:
        raise RuntimeError(f"wav must be float32 [C, L], got {wav.dtype} {tuple(wav.shape)}")


def _check_window(window: Optional[torch.Tensor], dev):
    if window is None:
        return
    if window.device != dev or window.dtype != torch.float32 or window.numel() != N_FFT or not window.is_contiguous():
        raise RuntimeError("window must be a contiguous float32 [2048] tensor on the input's device")


def stft_pack(wav: torch.Tensor, window: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[C,L] -> [C,2050,T] with row 2k = Re, 2k+1 = Im of bin k (torch.stft(n_fft=2048,
    hop=1024, center=True, reflect) + stack/reshape, m_dataset.py:187-190)."""
    _check_wav(wav)
    _check_window(window, wav.device)
    l = _lib.lib()
    wav = wav.contiguous()
    C, L = wav.shape
    T = l.bsb_num_frames(L)
    if T < 1:
        raise RuntimeError(f"wav too short for reflect padding: L={L} (need > 1024)")
    with torch.cuda.device(wav.device):
        x = torch.empty((C, 2050, T), dtype=torch.float32, device=wav.device)
        ws = torch.empty(C * T * 2052 * 4, dtype=torch.uint8, device=wav.device)
        _lib.check(l.bsb_stft_pack(wav.data_ptr(), C, L, _p(window), x.data_ptr(), ws.data_ptr(), ws.numel(),
                                   _stream_ptr()), "bsb_stft_pack")
    return x


def unpack_istft(x: torch.Tensor, window: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[C,2050,T] -> [C,1024*(T-1)] (reshape to complex + torch.istft, m_dataset.py:192-195)."""
    if not isinstance(x, torch.Tensor) or not x.is_cuda or x.dtype != torch.float32 or x.dim() != 3 or x.shape[1] != 2050:
        raise RuntimeError("x must be a float32 CUDA tensor [C, 2050, T]")
    _check_window(window, x.device)
    l = _lib.lib()
    x = x.contiguous()
    C, _, T = x.shape
    if T < 2:
        raise RuntimeError("istft needs at least 2 frames")
    with torch.cuda.device(x.device):
        wav = torch.empty((C, HOP * (T - 1)), dtype=torch.float32, device=x.device)
        ws = torch.empty(C * T * 2052 * 4, dtype=torch.uint8, device=x.device)
        _lib.check(l.bsb_unpack_istft(x.data_ptr(), C, T, _p(window), wav.data_ptr(), ws.data_ptr(), ws.numel(),
                                      _stream_ptr()), "bsb_unpack_istft")
    return wav


def infer(model: BSRNN, waveform: torch.Tensor, window: Optional[torch.Tensor] = None):
    """Forward-only part of m_dataset.infer (m_dataset.py:182-195): returns (x, x_time) with x the
    masked spectrogram [C,2050,T] and x_time the separated waveform."""
    spec = stft_pack(waveform, window)
    x = model.forward(spec)
    return x, unpack_istft(x, window)


def pair_stats(est: torch.Tensor, ref: torch.Tensor) -> torch.Tensor:
    """Per-row sums over an (estimate, reference) pair of float32 CUDA tensors [rows, ...] in ONE pass:
    returns float64 [rows, 5] = sum|e-r|, sum r^2, sum (e-r)^2, sum e*r, sum e^2 (bsb_pair_stats)."""
    for t in (est, ref):
        if not isinstance(t, torch.Tensor) or not t.is_cuda or t.dtype != torch.float32:
            raise RuntimeError("pair_stats takes float32 CUDA tensors: speechseparation_b200 has no CPU path")
    if est.shape != ref.shape or est.dim() < 2 or est.device != ref.device:
        raise RuntimeError(f"pair_stats: shapes/devices differ or no row dimension: {tuple(est.shape)} vs {tuple(ref.shape)}")
    rows = est.shape[0]
    e = est.reshape(rows, -1)
    r = ref.reshape(rows, -1)
    if e.stride(1) != 1:
        e = e.contiguous()
    if r.stride(1) != 1:
        r = r.contiguous()
    with torch.cuda.device(est.device):
        stats = torch.empty((rows, 5), dtype=torch.float64, device=est.device)
        _lib.check(_lib.lib().bsb_pair_stats(e.data_ptr(), e.stride(0) if rows > 1 else e.shape[1], r.data_ptr(),
                                             r.stride(0) if rows > 1 else r.shape[1], rows, e.shape[1],
                                             stats.data_ptr(), _stream_ptr()), "bsb_pair_stats")
    return stats


def metrics_from_stats(stats: torch.Tensor, n: int):
    """(l1_mean, sdr[rows], si_sdr[rows]) from pair_stats sums; n = elements per row.
    SDR as m_dataset.py:214-217 (+1e-9 on both energies); SI-SDR as the torchmetrics default the reference
    uses (m_dataset.py:201,224: zero_mean=False, eps = float32 eps)."""
    abs_d, rr, dd, er, ee = (stats[:, i] for i in range(5))
    l1 = abs_d.sum() / (stats.shape[0] * n)
    sdr = 10.0 * torch.log10((rr + 1e-9) / (dd + 1e-9))
    eps = torch.finfo(torch.float32).eps
    alpha = (er + eps) / (rr + eps)
    tgt = alpha * alpha * rr
    noise = (tgt - 2.0 * alpha * er + ee).clamp_min(0.0)
    si = 10.0 * torch.log10((tgt + eps) / (noise + eps))
    return l1, sdr, si


def train_infer(model: BSRNN, sample, window: Optional[torch.Tensor] = None):
    """Forward-only mirror of m_dataset.train_infer(model, None, sample, L1Loss) (m_dataset.py:202-226), the
    body of train.py's no_grad validation loop (train.py:132-152): returns (loss, sdr, sdr2, sdr3) as
    0-dim float64 CUDA tensors. sample = (mixture, speech) float32 CUDA tensors, either [1,C,L] as the reference's
    DataLoader (batch_size=1, train.py:84-85) hands them over, or [C,L].
    sdr2 follows the reference's reduction literally: it sums over `dim=1` of the UN-squeezed sample
    (m_dataset.py:219-222), which for [1,C,L] is the channel axis -- a mean over the L samples of a channel-summed
    log ratio (bsb_channel_sdr) -- and for a [C,L] sample the time axis (per-row SDR of the mixture, mean over rows).
    All reductions run on the device in one pass per pair (no torchmetrics)."""
    waveform, speech = sample
    batched = waveform.dim() == 3
    if batched:
        if waveform.shape[0] != 1 or speech.shape[0] != 1:
            raise RuntimeError("train_infer: the leading batch dimension must be 1 (m_dataset.py:184-185 squeezes it)")
        waveform, speech = waveform.squeeze(0), speech.squeeze(0)
    _check_wav(waveform)
    _check_wav(speech)
    x, x_time = infer(model, waveform, window)
    speech_freq = stft_pack(speech, window)
    speech_time = speech[:, :x_time.shape[1]]
    n_t = x_time.shape[1]
    st_time = pair_stats(x_time, speech_time)
    st_freq = pair_stats(x, speech_freq)
    st_mix = None if batched else pair_stats(waveform, speech)
    l1_time, sdr, sisdr = metrics_from_stats(st_time, n_t)
    # L1(Re) + L1(Im): each a mean over C*1025*T elements = the packed array's sum / (C*1025*T)
    l1_freq = st_freq[:, 0].sum() / (x.shape[0] * (x.shape[1] // 2) * x.shape[2])
    # sdr2 (m_dataset.py:219-222): n2Delta = |speech - mixture|^2: here the "estimate" is the mixture
    if batched:
        sdr2 = channel_sdr(waveform, speech)
    else:
        sdr2 = metrics_from_stats(st_mix, waveform.shape[1])[1].mean()
    return l1_time + l1_freq, sdr.mean(), sdr2, sisdr.mean()


def channel_sdr(mixture: torch.Tensor, speech: torch.Tensor) -> torch.Tensor:
    """mean over samples of 10 log10((sum_c s^2 + 1e-9) / (sum_c (s - m)^2 + 1e-9)) for [C,L] float32 CUDA tensors:
    the reference's sdr2 on its [1,C,L] samples (m_dataset.py:219-222). 0-dim float64 CUDA tensor."""
    _check_wav(mixture)
    _check_wav(speech)
    if mixture.shape != speech.shape or mixture.device != speech.device:
        raise RuntimeError("channel_sdr: shapes/devices differ")
    m, s_ = mixture.contiguous(), speech.contiguous()
    C, L = m.shape
    with torch.cuda.device(m.device):
        out = torch.empty((), dtype=torch.float64, device=m.device)
        _lib.check(_lib.lib().bsb_channel_sdr(m.data_ptr(), L, s_.data_ptr(), L, C, L, out.data_ptr(), _stream_ptr()),
                   "bsb_channel_sdr")
    return out / L


class StreamingSeparator:
    """Hop-by-hop separation with carried LSTM state (infer-streaming.py:84-147).

    All rolling state (2048-sample ring, LSTM state, overlap-add tail) stays on the device;
    ``step`` takes 1024 new samples per row and returns the 1024 output samples of that hop
    (the audio of the previous hop: one-hop algorithmic delay, as the reference).  After the
    first (warm-up) hop the whole hop is replayed as one CUDA graph.
    """

    def __init__(self, model: BSRNN, rows: int = 2, device=None, window: Optional[torch.Tensor] = None,
                 use_graph: bool = True):
        self.model = model
        self.rows = int(rows)
        dev = torch.device(device) if device is not None else next(model.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("StreamingSeparator needs the model on a CUDA device (no CPU path)")
        self.device = dev
        _check_window(window, dev)
        self.window = window
        l = _lib.lib()
        C = self.rows
        with torch.cuda.device(dev):
            self.ring = torch.zeros((C, N_FFT), dtype=torch.float32, device=dev)
            self.tail = torch.zeros((C, HOP), dtype=torch.float32, device=dev)
            self.state = [torch.zeros((4, 2, C * 12, 64), dtype=torch.float32, device=dev) for _ in range(2)]
            self.chunk = torch.zeros((C, HOP), dtype=torch.float32, device=dev)
            self.out = torch.zeros((C, HOP), dtype=torch.float32, device=dev)
            self.ws = torch.empty(l.bsb_stream_workspace_bytes(C), dtype=torch.uint8, device=dev)
        self.packed = model.packed_weights(dev)
        self.hops = 0
        self.use_graph = use_graph
        self._graphs = None

    def reset(self):
        for t in (self.ring, self.tail, self.state[0], self.state[1]):
            t.zero_()
        self.hops = 0

    def _enqueue(self, parity: int):
        l = _lib.lib()
        _lib.check(l.bsb_stream_step(self.chunk.data_ptr(), self.rows, _p(self.window), self.packed.data_ptr(),
                                     self.ring.data_ptr(), self.state[parity].data_ptr(),
                                     self.state[parity ^ 1].data_ptr(), self.tail.data_ptr(), self.out.data_ptr(),
                                     self.ws.data_ptr(), self.ws.numel(), _stream_ptr()), "bsb_stream_step")

    def _capture(self):
        # two graphs: the LSTM state ping-pongs between two buffers
        graphs = []
        saved = [t.clone() for t in (self.ring, self.tail, self.state[0], self.state[1])]
        for parity in (0, 1):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._enqueue(parity)
            graphs.append(g)
        for t, s in zip((self.ring, self.tail, self.state[0], self.state[1]), saved):
            t.copy_(s)          # capture does not execute, but keep the state provably untouched
        self._graphs = graphs

    @property
    def current_state(self) -> torch.Tensor:
        return self.state[self.hops & 1]

    def step(self, chunk: torch.Tensor) -> torch.Tensor:
        """chunk [rows,1024] float32 (CUDA, or pinned/pageable host memory: copied in) ->
        view of the device output buffer [rows,1024] (valid until the next step)."""
        if chunk.shape != self.chunk.shape or chunk.dtype != torch.float32:
            raise RuntimeError(f"chunk must be float32 {tuple(self.chunk.shape)}, got {chunk.dtype} {tuple(chunk.shape)}")
        with torch.cuda.device(self.device):
            self.chunk.copy_(chunk, non_blocking=True)
            parity = self.hops & 1
            if self.use_graph and self.hops >= 2:
                if self._graphs is None:
                    self._capture()
                self._graphs[parity].replay()
            else:
                self._enqueue(parity)
            self.hops += 1
        return self.out


class BatchSeparator:
    """Host-buffer front end of ``BSRNN.separate`` for bulk offline separation (the loop of
    gendataset-separate.py / infer.py over many clips): batches of [rows, L] arrive in pinned host
    memory and leave as [rows, 1024*(T-1)] in pinned host memory.

    The three legs of a batch -- host->device copy, the separation kernels, device->host copy -- run on
    three streams over ``depth`` device buffer pairs, so the copies of neighbouring batches overlap the
    kernels (PCIe is full duplex). Results are identical to calling ``model.separate`` per batch: the
    kernels of consecutive batches still run one after the other on one stream and share one workspace.
    """

    class Ticket:
        def __init__(self, event, out):
            self.event, self.out = event, out

        def wait(self) -> torch.Tensor:
            self.event.synchronize()
            return self.out

    def __init__(self, model: BSRNN, rows: int, samples: int, device=None, depth: int = 2,
                 window: Optional[torch.Tensor] = None, pcm16: bool = False):
        """pcm16=True: batches arrive and leave as int16 PCM (what the clips' wav files hold; x = s / 32768 as
        torchaudio.load normalises them, infer.py:24-27) -- half the host<->device bytes, converted on the device
        (bsb_pcm16_to_float / bsb_float_to_pcm16). float32 (default) is the reference's tensor contract."""
        dev = torch.device(device) if device is not None else next(model.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("BatchSeparator needs a CUDA device: speechseparation_b200 has no CPU path")
        T = _lib.lib().bsb_num_frames(samples)
        if T < 2 or rows < 1 or depth < 1:
            raise RuntimeError(f"bad batch shape rows={rows} samples={samples} depth={depth}")
        _check_window(window, dev)
        self.model, self.dev, self.rows, self.samples, self.depth, self.window = model, dev, rows, samples, depth, window
        self.out_samples = HOP * (T - 1)
        self.pcm16 = bool(pcm16)
        self.io_dtype = torch.int16 if self.pcm16 else torch.float32
        with torch.cuda.device(dev):
            self.s_in, self.s_run, self.s_out = (torch.cuda.Stream(dev) for _ in range(3))
            self.d_in = [torch.empty(rows, samples, device=dev, dtype=self.io_dtype) for _ in range(depth)]
            self.d_out = [torch.empty(rows, self.out_samples, device=dev, dtype=self.io_dtype) for _ in range(depth)]
            if self.pcm16:      # float staging of one batch (the kernels of consecutive batches run one after the other)
                self.f_in = torch.empty(rows, samples, device=dev)
                self.f_out = torch.empty(rows, self.out_samples, device=dev)
            self.ev_in = [torch.cuda.Event() for _ in range(depth)]
            self.ev_run = [torch.cuda.Event() for _ in range(depth)]
            self.ev_out = [torch.cuda.Event() for _ in range(depth)]
        self.used = [False] * depth
        self.n = 0

    def _check_host(self, t: torch.Tensor, shape, what: str):
        if t.device.type != "cpu" or t.dtype != self.io_dtype or tuple(t.shape) != tuple(shape) or not t.is_contiguous():
            raise RuntimeError(f"{what} must be a contiguous {self.io_dtype} host tensor of shape {tuple(shape)}")
        if not t.is_pinned():
            raise RuntimeError(f"{what} must be in pinned host memory (torch.empty(..., pin_memory=True))")

    def after(self, event: torch.cuda.Event):
        """Order all three legs after ``event`` (e.g. a timing event of the caller's stream)."""
        for s in (self.s_in, self.s_run, self.s_out):
            s.wait_event(event)

    def submit(self, host_in: torch.Tensor, host_out: torch.Tensor) -> "BatchSeparator.Ticket":
        """Enqueue one batch; returns at once. ``host_out`` is valid after ``ticket.wait()``."""
        self._check_host(host_in, (self.rows, self.samples), "host_in")
        self._check_host(host_out, (self.rows, self.out_samples), "host_out")
        b = self.n % self.depth
        self.n += 1
        with torch.cuda.device(self.dev):
            if self.used[b]:
                self.s_in.wait_event(self.ev_run[b])          # d_in[b] is free once its kernels are done
            with torch.cuda.stream(self.s_in):
                self.d_in[b].copy_(host_in, non_blocking=True)
                self.ev_in[b].record()
            self.s_run.wait_event(self.ev_in[b])
            if self.used[b]:
                self.s_run.wait_event(self.ev_out[b])         # d_out[b] is free once it reached the host
            with torch.cuda.stream(self.s_run):
                if self.pcm16:
                    l, sp = _lib.lib(), _stream_ptr()
                    _lib.check(l.bsb_pcm16_to_float(self.d_in[b].data_ptr(), self.d_in[b].numel(), self.f_in.data_ptr(), sp),
                               "bsb_pcm16_to_float")
                    self.model.separate(self.f_in, self.window, out=self.f_out)
                    _lib.check(l.bsb_float_to_pcm16(self.f_out.data_ptr(), self.f_out.numel(), self.d_out[b].data_ptr(), sp),
                               "bsb_float_to_pcm16")
                else:
                    self.model.separate(self.d_in[b], self.window, out=self.d_out[b])
                self.ev_run[b].record()
            self.s_out.wait_event(self.ev_run[b])
            with torch.cuda.stream(self.s_out):
                host_out.copy_(self.d_out[b], non_blocking=True)
                self.ev_out[b].record()
        self.used[b] = True
        return BatchSeparator.Ticket(self.ev_out[b], host_out)

    def join(self, stream: Optional[torch.cuda.Stream] = None):
        """Make ``stream`` (default: the current one) wait for everything submitted so far."""
        stream = stream if stream is not None else torch.cuda.current_stream(self.dev)
        for b in range(self.depth):
            if self.used[b]:
                stream.wait_event(self.ev_out[b])
