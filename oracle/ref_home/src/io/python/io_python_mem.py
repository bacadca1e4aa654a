"""In-memory IO back-end for the compiled reference (oracle/_ref/magritte/core.so).

TEST INFRASTRUCTURE ONLY. The reference's `IoPython("mem", name)` imports a module
`io_python_mem` from `MAGRITTE_FOLDER/src/io/python/` (reference:
src/io/python/io_python.cpp:11,18,449-519) and calls the nine functions below. The
reference ships an h5py back-end (src/io/python/io_python_hdf5.py); h5py is not
available in this image, so this module keeps the same nine-function contract over a
process-local dict. Semantics mirrored from io_python_hdf5.py:1-166:
  * "a/b/.x" is attribute x of object "a/b/" (empty object = file root);
  * read_length/read_width of a non-dataset count the children of the parent group
    whose name contains the last path component;
  * read_array of a missing dataset returns None.
"""
import numpy as np

# io_file -> {"data": {path: ndarray}, "attrs": {(obj, attr): value}}
FILES = {}


def _f(io_file, create=False):
    if create and io_file not in FILES:
        FILES[io_file] = {"data": {}, "attrs": {}}
    return FILES[io_file]


def _norm(path):
    return path.strip("/")


def file_exists(file_name):
    return file_name in FILES


def _children(f, group):
    group = _norm(group)
    pre = group + "/" if group else ""
    kids = set()
    for key in list(f["data"].keys()) + [k[0] for k in f["attrs"].keys()]:
        key = _norm(key)
        if key.startswith(pre) and len(key) > len(pre):
            kids.add(key[len(pre):].split("/")[0])
    return kids


def _count(f, file_name):
    object_name = file_name.split("/")[-1]
    group_name = file_name[: -len(object_name)] if object_name else file_name
    kids = _children(f, group_name)
    if not kids and _norm(group_name):
        raise KeyError(group_name)
    return sum(1 for k in kids if object_name in k)


def read_length(io_file, file_name):
    f = _f(io_file)
    key = _norm(file_name)
    if key in f["data"]:
        return len(f["data"][key])
    return _count(f, file_name)


def read_width(io_file, file_name):
    f = _f(io_file)
    key = _norm(file_name)
    if key in f["data"]:
        return f["data"][key].shape[1]
    return _count(f, file_name)


def _split(file_name):
    parts = file_name.split(".")
    return _norm(parts[0]), parts[1]


def read_attribute(io_file, file_name):
    f = _f(io_file)
    return f["attrs"][_split(file_name)]


def write_attribute(io_file, file_name, data):
    f = _f(io_file, create=True)
    f["attrs"][_split(file_name)] = data


def read_number(io_file, file_name):
    return read_attribute(io_file, file_name)


def write_number(io_file, file_name, data):
    write_attribute(io_file, file_name, data)


def _element_type(l):
    if isinstance(l, list):
        return _element_type(l[0])
    return type(l)


def read_array(io_file, file_name):
    f = _f(io_file)
    key = _norm(file_name)
    if key in f["data"]:
        return np.array(f["data"][key])
    return None


def write_array(io_file, file_name, data):
    f = _f(io_file, create=True)
    if _element_type(data) == str:
        f["data"][_norm(file_name)] = np.array(data, dtype="S")
    else:
        f["data"][_norm(file_name)] = np.array(data)
