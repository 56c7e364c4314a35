"""Hand-assembled Keras-2.2.4 / h5py-2.x style HDF5 files, written from the HDF5 file-format specification and NOT with
cnn_itmo_b200.h5min.Writer, so that h5min's reader (the code that has to open the real saved7-model-218-0.73.hdf5 of
/root/reference/predict.py:24 on day one) is exercised on structures its own writer never emits:

  * superblock v0 with non-default B-tree K values; symbol-table groups whose v1 B-trees are SEVERAL LEVELS deep
    (internal nodes over leaf nodes over many SNODs), entries in strcmp order — NOT creation / layer_names order
  * v1 object headers whose attribute messages sit in CONTINUATION blocks (what libhdf5 does when attributes are added
    after the object was created, as Keras does), with NIL padding messages and message types a reader must skip
    (modification time 0x0012, fill value 0x0005)
  * attributes as h5py writes them for Keras: FIXED-LENGTH null-padded byte strings for bytes objects
    (keras_version, backend, model_config, training_config), fixed-length string ARRAYS (layer_names, weight_names),
    a zero-length float64 array for weightless layers, and a VARIABLE-LENGTH string through a global heap collection
  * datasets in CONTIGUOUS layout (h5py's default, what Keras produces) and in CHUNKED layout with a v1 chunk B-tree
    whose edge chunks overhang the dataset (uncompressed), little-endian IEEE float32

`build_keras_file(weights, ...)` returns the file's bytes for the full 19-layer U-Net; tests/test_h5min_cpu.py writes them
to a temporary file and reads them back through cnn_itmo_b200.load_model's reader. Run as a script to write a file.
"""
import json
import struct
import sys

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF


def _pad8(b: bytes) -> bytes:
    return b + b'\x00' * (-len(b) % 8)


class Assembler:
    def __init__(self, leaf_k=1, internal_k=2, chunked_every=0, attrs_in_continuation=True):
        self.buf = bytearray(b'\x00' * 96)           # superblock v0 (56 bytes + 40-byte root symbol-table entry)
        self.leaf_k, self.internal_k = leaf_k, internal_k
        self.chunked_every = chunked_every
        self.cont = attrs_in_continuation
        self.n_datasets = 0
        self.gheap = []                               # (index, bytes) objects of ONE global heap collection
        self.gheap_addr_fixups = []                   # positions of 8-byte collection addresses to patch

    # ---- raw allocation ------------------------------------------------------------------------------------------
    def alloc(self, data: bytes, align=8) -> int:
        while len(self.buf) % align:
            self.buf.append(0)
        addr = len(self.buf)
        self.buf += data
        return addr

    # ---- datatypes / dataspaces ------------------------------------------------------------------------------------
    @staticmethod
    def dt_f32() -> bytes:   # class 1 (float), version 1; little-endian, IEEE single
        return struct.pack('<BBBBI', 0x11, 0x20, 0x1F, 0x00, 4) + struct.pack('<HHBBBBI', 0, 32, 23, 8, 0, 23, 127)

    @staticmethod
    def dt_f64() -> bytes:
        return struct.pack('<BBBBI', 0x11, 0x20, 0x3F, 0x00, 8) + struct.pack('<HHBBBBI', 0, 64, 52, 11, 0, 52, 1023)

    @staticmethod
    def dt_fixed_str(n: int) -> bytes:   # class 3 (string), version 1; null-PADDED (numpy 'S'), ASCII
        return struct.pack('<BBBBI', 0x13, 0x01, 0x00, 0x00, n)

    @staticmethod
    def dt_vlen_str() -> bytes:          # class 9 (variable length), type string, null-terminated, UTF-8; base = 1-byte string
        return struct.pack('<BBBBI', 0x19, 0x01, 0x01, 0x00, 16) + struct.pack('<BBBBI', 0x13, 0x00, 0x00, 0x00, 1)

    @staticmethod
    def ds(shape) -> bytes:              # dataspace message version 1
        return struct.pack('<BBB5x', 1, len(shape), 0) + b''.join(struct.pack('<Q', d) for d in shape)

    # ---- attribute messages (version 1) ----------------------------------------------------------------------------
    def attr(self, name: str, value) -> bytes:
        nm = name.encode() + b'\x00'
        if isinstance(value, tuple) and value[0] == 'vlen':          # variable-length string scalar via the global heap
            raw = value[1]
            idx = len(self.gheap) + 1
            self.gheap.append((idx, raw))
            dt, sp = self.dt_vlen_str(), self.ds(())
            data = struct.pack('<I', len(raw)) + b'\x00' * 8 + struct.pack('<I', idx)
            fix = ('gheap', len(data) - 12)
        elif isinstance(value, bytes):                                 # h5py: bytes -> fixed-length string scalar
            dt, sp, data, fix = self.dt_fixed_str(max(1, len(value))), self.ds(()), value or b'\x00', None
        else:
            a = np.asarray(value)
            if a.dtype.kind == 'S':
                dt = self.dt_fixed_str(a.dtype.itemsize)
            elif a.dtype == np.float64:
                dt = self.dt_f64()
            else:
                raise TypeError(a.dtype)
            sp, data, fix = self.ds(a.shape), a.tobytes(), None
        head = struct.pack('<BBHHH', 1, 0, len(nm), len(dt), len(sp))
        body = head + _pad8(nm) + _pad8(dt) + _pad8(sp)
        off = len(body)
        body += data
        return body, ((off + fix[1]) if fix else None)

    # ---- object headers (version 1) --------------------------------------------------------------------------------
    def object_header(self, first_msgs, attr_msgs) -> int:
        """first_msgs: [(type, body)] placed in the header's own block; attr_msgs: [(body, gheap_fixup)] placed in a
        continuation block (or in the first block when continuation is off)."""
        def enc(msgs):
            out = bytearray()
            fixups = []
            for mtype, body, fix in msgs:
                b = _pad8(body)
                out += struct.pack('<HHB3x', mtype, len(b), 0)
                if fix is not None:
                    fixups.append(len(out) + fix)
                out += b
            return bytes(out), fixups
        msgs = [(t, b, None) for t, b in first_msgs]
        amsgs = [(0x000C, b, f) for b, f in attr_msgs]
        if not self.cont or not amsgs:
            block, fixups = enc(msgs + amsgs)
            hdr = struct.pack('<BBHII4x', 1, 0, len(msgs) + len(amsgs), 1, len(block))
            addr = self.alloc(hdr + block)
            for f in fixups:
                self.gheap_addr_fixups.append(addr + 16 + f)
            return addr
        # continuation block first (its address goes into the header), with a NIL message as padding in front
        nil = (0x0000, b'\x00' * 8, None)
        cblock, cfix = enc([nil] + amsgs)
        caddr = self.alloc(cblock)
        for f in cfix:
            self.gheap_addr_fixups.append(caddr + f)
        cont = (0x0010, struct.pack('<QQ', caddr, len(cblock)), None)
        block, _ = enc(msgs + [cont])
        hdr = struct.pack('<BBHII4x', 1, 0, len(msgs) + 1 + 1 + len(amsgs), 1, len(block))
        return self.alloc(hdr + block)

    # ---- datasets --------------------------------------------------------------------------------------------------
    def dataset(self, arr: np.ndarray) -> int:
        arr = np.ascontiguousarray(arr, dtype='<f4')
        self.n_datasets += 1
        chunked = self.chunked_every and self.n_datasets % self.chunked_every == 0 and arr.ndim >= 1 and arr.size > 0
        mtime = (0x0012, struct.pack('<B3xI', 1, 1546300800))
        fill = (0x0005, struct.pack('<BBBB', 2, 2, 0, 0))             # fill value v2: late allocation... undefined value
        if not chunked:
            daddr = self.alloc(arr.tobytes()) if arr.size else UNDEF
            layout = struct.pack('<BB', 3, 1) + struct.pack('<QQ', daddr, arr.nbytes)
        else:
            cdims = tuple(max(1, (d + 1) // 2 + (1 if d > 3 else 0)) for d in arr.shape)   # edge chunks overhang
            grid = [range(0, d, c) for d, c in zip(arr.shape, cdims)]
            entries = []
            for idx in np.ndindex(*[len(g) for g in grid]):
                origin = tuple(g[i] for g, i in zip(grid, idx))
                chunk = np.zeros(cdims, '<f4')
                sl = tuple(slice(o, min(o + c, d)) for o, c, d in zip(origin, cdims, arr.shape))
                view = arr[sl]
                chunk[tuple(slice(0, s) for s in view.shape)] = view
                entries.append((origin, self.alloc(chunk.tobytes())))
            csize = int(np.prod(cdims)) * 4
            rank1 = arr.ndim + 1

            def key(origin):
                return struct.pack('<II', csize, 0) + b''.join(struct.pack('<Q', o) for o in origin) + struct.pack('<Q', 0)
            node = bytearray(b'TREE' + struct.pack('<BBHQQ', 1, 0, len(entries), UNDEF, UNDEF))
            for origin, caddr in entries:
                node += key(origin) + struct.pack('<Q', caddr)
            node += key(tuple(arr.shape))                                  # final key: one past the last chunk
            baddr = self.alloc(bytes(node))
            layout = struct.pack('<BBB', 3, 2, rank1) + struct.pack('<Q', baddr) + \
                b''.join(struct.pack('<I', c) for c in cdims) + struct.pack('<I', 4)
        msgs = [(0x0001, self.ds(arr.shape)), (0x0003, self.dt_f32()), fill, (0x0008, layout), mtime]
        return self.object_header(msgs, [])

    # ---- groups (symbol table: local heap + v1 B-tree over SNODs) ---------------------------------------------------
    def group(self, children: dict, attrs: dict):
        """children: name -> object header address. Returns (object header address, btree address, heap address)."""
        names = sorted(children, key=lambda s: s.encode())
        heap = bytearray(b'\x00' * 8)                                     # offset 0: the empty string
        off = {}
        for n in names:
            off[n] = len(heap)
            heap += _pad8(n.encode() + b'\x00')
        free_off = len(heap)
        heap += struct.pack('<QQ', 1, 32) + b'\x00' * 16                  # one free block at the end
        haddr_data = self.alloc(bytes(heap))
        haddr = self.alloc(b'HEAP' + struct.pack('<B3xQQQ', 0, len(heap), free_off, haddr_data))
        per = 2 * self.leaf_k
        leaves = []                                                       # (address, last name's heap offset)
        for i in range(0, max(1, len(names)), per):
            part = names[i:i + per]
            body = bytearray(b'SNOD' + struct.pack('<BBH', 1, 0, len(part)))
            for n in part:
                body += struct.pack('<QQII16x', off[n], children[n], 0, 0)
            body += b'\x00' * (40 * (per - len(part)))
            leaves.append((self.alloc(bytes(body)), off[part[-1]] if part else 0))
        level, nodes = 0, leaves
        fan = 2 * self.internal_k
        while True:
            groups = [nodes[i:i + fan] for i in range(0, len(nodes), fan)]
            new = []
            addrs = []
            for g in groups:
                body = bytearray(b'TREE' + struct.pack('<BBH', 0, level, len(g)) + b'\x00' * 16)
                body += struct.pack('<Q', 0)                              # key 0: the empty string
                for a, last in g:
                    body += struct.pack('<QQ', a, last)
                body += b'\x00' * (16 * (fan - len(g)))
                addrs.append(self.alloc(bytes(body)))
                new.append((addrs[-1], g[-1][1]))
            for k, a in enumerate(addrs):                                 # sibling pointers
                left = addrs[k - 1] if k else UNDEF
                right = addrs[k + 1] if k + 1 < len(addrs) else UNDEF
                self.buf[a + 8:a + 24] = struct.pack('<QQ', left, right)
            nodes, level = new, level + 1
            if len(nodes) == 1:
                break
        baddr = nodes[0][0]
        stab = (0x0011, struct.pack('<QQ', baddr, haddr))
        ohdr = self.object_header([stab], [self.attr(k, v) for k, v in attrs.items()])
        return ohdr, baddr, haddr

    def finish(self, root_children: dict, root_attrs: dict) -> bytes:
        ohdr, baddr, haddr = self.group(root_children, root_attrs)
        if self.gheap:
            body = bytearray()
            for idx, raw in self.gheap:
                body += struct.pack('<HH4xQ', idx, 1, len(raw)) + _pad8(raw)
            free = 4096 - ((16 + len(body)) % 4096)
            if free < 16:
                free += 4096
            body += struct.pack('<HH4xQ', 0, 0, free - 16) + b'\x00' * (free - 16)
            coll = b'GCOL' + struct.pack('<B3xQ', 1, 16 + len(body)) + bytes(body)
            gaddr = self.alloc(coll)
            for pos in self.gheap_addr_fixups:
                self.buf[pos:pos + 8] = struct.pack('<Q', gaddr)
        eof = len(self.buf)
        sb = b'\x89HDF\r\n\x1a\n' + struct.pack('<BBBBBBBB', 0, 0, 0, 0, 0, 8, 8, 0) + \
            struct.pack('<HHI', self.leaf_k, self.internal_k, 0) + struct.pack('<QQQQ', 0, UNDEF, eof, UNDEF) + \
            struct.pack('<QQII', 0, ohdr, 1, 0) + struct.pack('<QQ', baddr, haddr)
        assert len(sb) == 96
        self.buf[0:96] = sb
        return bytes(self.buf)


def build_keras_file(weights, keras_layers, model_config, training_config=None, **kw) -> bytes:
    """weights: the repo's per-layer dicts in execution order (kernel, bias, gamma, beta, moving_mean, moving_variance);
    keras_layers: cnn_itmo_b200.graph.keras_layers() (all 66 Keras layers, model order)."""
    a = Assembler(**kw)
    layer_groups = {}
    bn_of = {}                       # BatchNormalization layer name -> weight layer index
    for k in keras_layers:
        if k.cls == 'BatchNormalization':
            bn_of[k.name] = k.weight_layer
    for k in keras_layers:
        if k.cls in ('Conv2D', 'Conv2DTranspose'):
            w = weights[k.weight_layer]
            items = [('kernel:0', w['kernel']), ('bias:0', w['bias'])]
        elif k.cls == 'BatchNormalization':
            w = weights[bn_of[k.name]]
            items = [('gamma:0', w['gamma']), ('beta:0', w['beta']), ('moving_mean:0', w['moving_mean']),
                     ('moving_variance:0', w['moving_variance'])]
        else:
            items = []
        if items:
            inner = {n: a.dataset(v) for n, v in items}
            inner_grp, _, _ = a.group(inner, {})
            names = np.array([f'{k.name}/{n}'.encode() for n, _ in items])
            grp, _, _ = a.group({k.name: inner_grp}, {'weight_names': names})
        else:
            grp, _, _ = a.group({}, {'weight_names': np.zeros((0,), np.float64)})
        layer_groups[k.name] = grp
    top_attrs = {'layer_names': np.array([k.name.encode() for k in keras_layers]), 'backend': b'tensorflow',
                 'keras_version': b'2.2.4'}
    mw, _, _ = a.group(layer_groups, top_attrs)
    root_attrs = {'keras_version': b'2.2.4', 'backend': ('vlen', b'tensorflow'),
                  'model_config': json.dumps(model_config).encode()}
    if training_config is not None:
        root_attrs['training_config'] = json.dumps(training_config).encode()
    return a.finish({'model_weights': mw}, root_attrs)


if __name__ == '__main__':
    import os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    from cnn_itmo_b200 import checkpoint, graph
    from oracle import unet_ref as R
    out = sys.argv[1] if len(sys.argv) > 1 else 'h5py_style.hdf5'
    data = build_keras_file(R.synthetic_weights(1), graph.keras_layers(), checkpoint.model_config(),
                            checkpoint.training_config(), chunked_every=3)
    open(out, 'wb').write(data)
    print(out, len(data), 'bytes')
