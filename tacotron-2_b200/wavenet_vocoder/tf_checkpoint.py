"""TensorFlow checkpoint (tensor bundle, "V2" format) -> weight table, without TensorFlow  (SURVEY.md 8f-1).

The reference restores the WaveNet's exponential-moving-average shadow variables with a `tf.train.Saver`
(wavenet_vocoder/train.py:67-86, synthesizer.py:27-44).  A V2 checkpoint `<prefix>` is two kinds of files:

  <prefix>.index                 a LevelDB-style sorted string table: key "" -> BundleHeaderProto, every other key
                                 (a variable name) -> BundleEntryProto {dtype, shape, shard_id, offset, size, crc32c}
  <prefix>.data-000NN-of-000MM   the raw little-endian tensor bytes

`read_bundle` parses those from the published formats (table: blocks of prefix-compressed entries with restart points,
a 48-byte footer with the metaindex / index block handles and the magic 0xdb4775248b80fb57; protos: plain varint /
length-delimited wire format).  `load_tf_checkpoint` then maps the reference's variable names to the logical names of
weights.weight_specs (scope prefixes and the `/ExponentialMovingAverage` suffix are dropped; layers are found by their
`ResidualConv1DGLU_<l>` / `residual_block_*` / `input_convolution` / `final_convolution_<i>` /
`local_conditioning_upsampling_<i>` / `gc_embedding` components) and validates every shape.

STATUS: the file-format reader is checked against bundles written by `write_bundle` below (same specification, so it
pins the reader against the writer, not against TensorFlow) - no TensorFlow checkpoint exists in this environment, and
the exact Keras / variable_scope auto-naming of the reference's variables could not be confirmed (SURVEY.md appendix
A).  `load_tf_checkpoint` therefore reports every name it could not place instead of guessing.  Block compression
(snappy) is not supported: TensorFlow's BundleWriter writes its index uncompressed.
"""
import os
import re
import struct
from collections import OrderedDict

import numpy as np

TABLE_MAGIC = 0xdb4775248b80fb57
_DTYPES = {1: np.float32, 2: np.float64, 3: np.int32, 9: np.int64, 19: np.float16}   # tensorflow/core/framework/types.proto


def _varint(buf, pos):
	out, shift = 0, 0
	while True:
		b = buf[pos]
		pos += 1
		out |= (b & 0x7f) << shift
		if not b & 0x80:
			return out, pos
		shift += 7


def _put_varint(v):
	out = bytearray()
	while True:
		b = v & 0x7f
		v >>= 7
		if v:
			out.append(b | 0x80)
		else:
			out.append(b)
			return bytes(out)


def _proto_fields(buf):
	"""(field number, wire type, value) of a serialized message; value = int (varint, fixed) or bytes."""
	pos, n = 0, len(buf)
	while pos < n:
		key, pos = _varint(buf, pos)
		field, wt = key >> 3, key & 7
		if wt == 0:
			v, pos = _varint(buf, pos)
		elif wt == 1:
			v = struct.unpack_from('<Q', buf, pos)[0]
			pos += 8
		elif wt == 2:
			ln, pos = _varint(buf, pos)
			v = bytes(buf[pos:pos + ln])
			pos += ln
		elif wt == 5:
			v = struct.unpack_from('<I', buf, pos)[0]
			pos += 4
		else:
			raise ValueError('unsupported protobuf wire type {}'.format(wt))
		yield field, wt, v


def _block_handle(buf, pos):
	off, pos = _varint(buf, pos)
	size, pos = _varint(buf, pos)
	return (off, size), pos


def _read_block(data, handle):
	off, size = handle
	if data[off + size] != 0:
		raise ValueError('compressed table block (type {}): not supported'.format(data[off + size]))
	return data[off:off + size]


def _block_entries(block):
	"""(key, value) pairs of one table block (prefix-compressed keys; the restart array at the end is not needed for a scan)."""
	n_restarts = struct.unpack_from('<I', block, len(block) - 4)[0]
	end = len(block) - 4 - 4 * n_restarts
	pos, key = 0, b''
	while pos < end:
		shared, pos = _varint(block, pos)
		non_shared, pos = _varint(block, pos)
		vlen, pos = _varint(block, pos)
		key = key[:shared] + bytes(block[pos:pos + non_shared])
		pos += non_shared
		yield key, bytes(block[pos:pos + vlen])
		pos += vlen


def read_index(index_path):
	"""`<prefix>.index` -> (header dict, OrderedDict name -> entry dict with dtype, shape, shard_id, offset, size)."""
	data = open(index_path, 'rb').read()
	if len(data) < 48 or struct.unpack_from('<Q', data, len(data) - 8)[0] != TABLE_MAGIC:
		raise ValueError('{} is not a TensorFlow checkpoint index (bad table magic)'.format(index_path))
	footer = data[len(data) - 48:]
	_, pos = _block_handle(footer, 0)          # metaindex block: unused
	index_handle, _ = _block_handle(footer, pos)
	header, entries = {}, OrderedDict()
	for _, handle_bytes in _block_entries(_read_block(data, index_handle)):
		handle, _ = _block_handle(handle_bytes, 0)
		for key, value in _block_entries(_read_block(data, handle)):
			if key == b'':
				for f, _, v in _proto_fields(value):   # BundleHeaderProto: num_shards = 1, endianness = 2
					if f == 1:
						header['num_shards'] = v
					elif f == 2:
						header['endianness'] = v
				continue
			e = {'dtype': 0, 'shape': [], 'shard_id': 0, 'offset': 0, 'size': 0, 'sliced': False}
			for f, _, v in _proto_fields(value):       # BundleEntryProto
				if f == 1:
					e['dtype'] = v
				elif f == 2:                            # TensorShapeProto: repeated Dim dim = 2 {int64 size = 1}
					for f2, _, v2 in _proto_fields(v):
						if f2 == 2:
							size = 0
							for f3, _, v3 in _proto_fields(v2):
								if f3 == 1:
									size = v3
							e['shape'].append(size)
				elif f == 3:
					e['shard_id'] = v
				elif f == 4:
					e['offset'] = v
				elif f == 5:
					e['size'] = v
				elif f == 7:
					e['sliced'] = True
			entries[key.decode('utf-8')] = e
	header.setdefault('num_shards', 1)
	if header.get('endianness', 0) != 0:
		raise ValueError('big-endian checkpoint: not supported')
	return header, entries


def read_bundle(prefix, names=None):
	"""All (or the named) tensors of the checkpoint `<prefix>` as an OrderedDict name -> numpy array."""
	header, entries = read_index(prefix + '.index')
	shards = {}
	out = OrderedDict()
	for name, e in entries.items():
		if names is not None and name not in names:
			continue
		if e['sliced']:
			raise ValueError('variable {} is stored in slices (partitioned variable): not supported'.format(name))
		if e['dtype'] not in _DTYPES:
			continue                                     # strings / resources: nothing the vocoder needs
		sid = e['shard_id']
		if sid not in shards:
			shards[sid] = np.memmap('{}.data-{:05d}-of-{:05d}'.format(prefix, sid, header['num_shards']), dtype=np.uint8, mode='r')
		dt = np.dtype(_DTYPES[e['dtype']])
		count = int(np.prod(e['shape'])) if e['shape'] else 1
		if count * dt.itemsize != e['size']:
			raise ValueError('variable {}: {} bytes stored, shape {} needs {}'.format(name, e['size'], e['shape'], count * dt.itemsize))
		raw = shards[sid][e['offset']:e['offset'] + e['size']]
		out[name] = np.frombuffer(raw.tobytes(), dtype=dt).reshape(e['shape'])
	return out


# ---- name mapping -------------------------------------------------------------------------------------------------
_ROLE = re.compile(r'(causal_conv|cin_conv|gin_conv|out_conv|skip_conv)')


def logical_name(tf_name):
	"""Reference variable name -> name of weights.weight_specs, or None.  Works on the name's components, whatever
	scopes surround them: `.../ResidualConv1DGLU_3/residual_block_causal_conv_ResidualConv1DGLU_3/kernel/ExponentialMovingAverage`
	-> `ResidualConv1DGLU_3/residual_block_causal_conv/kernel`."""
	parts = [p for p in tf_name.split('/') if p not in ('ExponentialMovingAverage',)]
	if not parts:
		return None
	leaf = parts[-1].split(':')[0]
	if leaf in ('kernel', 'bias'):
		body = '/'.join(parts[:-1])
	elif 'embedding' in leaf or leaf == 'gc_embedding':
		return 'gc_embedding' if ('gc_embedding' in tf_name or 'embedding' in leaf) else None
	else:
		return None
	m = re.search(r'ResidualConv1DGLU_(\d+)', body)
	if m:
		role = _ROLE.search(body[m.start():])
		return 'ResidualConv1DGLU_{}/residual_block_{}/{}'.format(m.group(1), role.group(1), leaf) if role else None
	m = re.search(r'final_convolution_(\d+)', body)
	if m:
		return 'final_convolution_{}/{}'.format(m.group(1), leaf)
	if 'input_convolution' in body:
		return 'input_convolution/' + leaf
	m = re.search(r'local_conditioning_upsampling_(\d+)', body)
	if m:
		return 'local_conditioning_upsampling_{}/{}'.format(m.group(1), leaf)
	return None


def load_tf_checkpoint(prefix, hparams):
	"""Weight table (weights.weight_specs names, float32) from the TensorFlow checkpoint `<prefix>`.  Where a variable
	exists both raw and as its moving average the average wins (the reference synthesises from the EMA weights)."""
	from wavenet_vocoder import weights as wn_weights
	specs = wn_weights.weight_specs(hparams)
	tensors = read_bundle(prefix)
	table, from_ema, unplaced = OrderedDict(), set(), []
	for name, arr in tensors.items():
		if 'Adam' in name or 'global_step' in name or 'beta1_power' in name or 'beta2_power' in name:
			continue
		ln = logical_name(name)
		if ln is None or ln not in specs:
			unplaced.append(name)
			continue
		ema = name.endswith('ExponentialMovingAverage')
		if ln in table and ln in from_ema and not ema:
			continue
		a = np.asarray(arr, np.float32)
		want = tuple(specs[ln])
		if a.shape != want and a.size == int(np.prod(want)):
			a = a.reshape(want)                          # e.g. a Conv1D kernel stored with an extra unit axis
		if a.shape != want:
			raise ValueError('checkpoint variable {} has shape {}, {} needs {}'.format(name, arr.shape, ln, want))
		table[ln] = a
		if ema:
			from_ema.add(ln)
	missing = [n for n in specs if n not in table]
	if missing:
		raise RuntimeError('Failed to load checkpoint at {}: no variable found for {} (of {}); variables that could not be placed: {}'.format(
			prefix, missing[:6], len(missing), unplaced[:6]))
	return OrderedDict((n, table[n]) for n in specs)


# ---- writer (fixtures for the tests of the reader; block checksums are written as zero, which TensorFlow would reject) ----
def _build_block(items, restart_interval=16):
	out, restarts, prev = bytearray(), [], b''
	for i, (key, value) in enumerate(items):
		shared = 0
		if i % restart_interval == 0:
			restarts.append(len(out))
		else:
			while shared < min(len(prev), len(key)) and prev[shared] == key[shared]:
				shared += 1
		out += _put_varint(shared) + _put_varint(len(key) - shared) + _put_varint(len(value)) + key[shared:] + value
		prev = key
	if not restarts:
		restarts = [0]
	for r in restarts:
		out += struct.pack('<I', r)
	out += struct.pack('<I', len(restarts))
	return bytes(out)


def _field(num, wt, payload):
	return _put_varint((num << 3) | wt) + payload


def write_bundle(prefix, tensors, block_entries=7):
	"""Writes `<prefix>.index` and `<prefix>.data-00000-of-00001` in the tensor-bundle layout read_bundle parses."""
	inv = {np.dtype(v): k for k, v in _DTYPES.items()}
	data, entries = bytearray(), []
	for name in sorted(tensors):
		a = np.asarray(tensors[name])
		a = a if a.flags['C_CONTIGUOUS'] else np.ascontiguousarray(a)   # (ascontiguousarray alone turns a scalar into shape (1,))
		shape = b''.join(_field(2, 2, _put_varint(len(_field(1, 0, _put_varint(int(d))))) + _field(1, 0, _put_varint(int(d)))) for d in a.shape)
		e = _field(1, 0, _put_varint(inv[a.dtype])) + _field(2, 2, _put_varint(len(shape)) + shape) + \
			_field(4, 0, _put_varint(len(data))) + _field(5, 0, _put_varint(a.nbytes)) + _field(6, 5, struct.pack('<I', 0))
		entries.append((name.encode('utf-8'), e))
		data += a.tobytes()
	header = _field(1, 0, _put_varint(1)) + _field(2, 0, _put_varint(0))
	items = [(b'', header)] + entries
	out, index_items = bytearray(), []
	for i in range(0, len(items), block_entries):
		blk = _build_block(items[i:i + block_entries], restart_interval=3)
		index_items.append((items[min(i + block_entries, len(items)) - 1][0], _put_varint(len(out)) + _put_varint(len(blk))))
		out += blk + b'\x00' + struct.pack('<I', 0)
	meta = _build_block([])
	meta_handle = _put_varint(len(out)) + _put_varint(len(meta))
	out += meta + b'\x00' + struct.pack('<I', 0)
	idx = _build_block(index_items, restart_interval=1)
	idx_handle = _put_varint(len(out)) + _put_varint(len(idx))
	out += idx + b'\x00' + struct.pack('<I', 0)
	footer = meta_handle + idx_handle
	out += footer + b'\x00' * (40 - len(footer)) + struct.pack('<Q', TABLE_MAGIC)
	with open(prefix + '.index', 'wb') as f:
		f.write(bytes(out))
	with open(prefix + '.data-00000-of-00001', 'wb') as f:
		f.write(bytes(data))
