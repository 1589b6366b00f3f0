// packer.cpp -- _b200mt_pack: turns a Python list of lists of torch tensors into the raw
// (pointer, numel, dtype) arrays of the C ABI at C speed.  This is the only torch-aware native code in
// the tree and it contains no kernel: it exists because walking 4 x 400 tensors in Python costs more
// than the fused kernel that follows (the reference pays the same conversion inside pybind11,
// csrc/amp_C_frontend.cpp).  The checks and messages are those of csrc/multi_tensor_apply.cuh:53-70.
#include <Python.h>
#include <torch/csrc/autograd/python_variable.h>
#include <ATen/ATen.h>
#include <stdint.h>

namespace {

int dtype_code(at::ScalarType t) {
  switch (t) {
    case at::kFloat: return 0;
    case at::kHalf: return 1;
    case at::kBFloat16: return 2;
    case at::kDouble: return 3;
    case at::kByte: return 4;
    default: return -1;
  }
}

PyObject* fail(const char* msg) {
  PyErr_SetString(PyExc_RuntimeError, msg);
  return nullptr;
}

// pack_lists(tensor_lists, ptrs_addr, numel_addr, capacity_tensors)
//   -> (ntensors, depth, (dtype codes), device_index)   or None when an item is not a torch tensor
PyObject* pack_lists(PyObject*, PyObject* args) {
  PyObject* lists;
  unsigned long long ptrs_addr, numel_addr;
  long long capacity;
  if (!PyArg_ParseTuple(args, "OKKL", &lists, &ptrs_addr, &numel_addr, &capacity)) return nullptr;
  uint64_t* ptrs = reinterpret_cast<uint64_t*>(ptrs_addr);
  int64_t* numel = reinterpret_cast<int64_t*>(numel_addr);
  PyObject* outer = PySequence_Fast(lists, "tensor_lists must be a sequence of sequences");
  if (!outer) return nullptr;
  const Py_ssize_t depth = PySequence_Fast_GET_SIZE(outer);
  if (depth < 1 || depth > 5) { Py_DECREF(outer); return fail("tensor_lists.size() != depth"); }
  Py_ssize_t T = -1;
  int codes[5] = {0, 0, 0, 0, 0};
  c10::Device ref_device(c10::kCPU);
  for (Py_ssize_t d = 0; d < depth; d++) {
    PyObject* inner = PySequence_Fast(PySequence_Fast_GET_ITEM(outer, d), "tensor_lists must be a sequence of sequences");
    if (!inner) { Py_DECREF(outer); return nullptr; }
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(inner);
    const char* err = nullptr;
    bool not_tensor = false;
    if (d == 0) {
      T = n;
      if (T <= 0) err = "tensor_lists[0].size() is not > 0";
      else if ((long long)T * depth > capacity) err = "internal: pack buffer too small";
    } else if (n != T) {
      err = "Size mismatch among tensor lists";
    }
    at::ScalarType list_dtype = at::kFloat;
    for (Py_ssize_t t = 0; !err && t < n; t++) {
      PyObject* o = PySequence_Fast_GET_ITEM(inner, t);
      if (!THPVariable_Check(o)) { not_tensor = true; break; }
      const at::Tensor& x = THPVariable_Unpack(o);
      if (d == 0 && t == 0) {
        ref_device = x.device();
        if (!ref_device.is_cuda()) { err = "expected input to be on cuda"; break; }
      }
      if (t == 0) {
        list_dtype = x.scalar_type();
        codes[d] = dtype_code(list_dtype);
      }
      if (x.is_sparse()) { err = "sparse tensors are not supported on this path"; break; }
      bool contiguous = x.is_contiguous();
      if (!contiguous && x.dim() == 4) contiguous = x.is_contiguous(at::MemoryFormat::ChannelsLast);
      if (!contiguous && x.dim() == 5) contiguous = x.is_contiguous(at::MemoryFormat::ChannelsLast3d);
      if (!contiguous) { err = "A tensor was not contiguous."; break; }
      if (x.device() != ref_device) { err = "A tensor was not on the same device as the first tensor"; break; }
      const int64_t ne = x.numel();
      if (d == 0) numel[t] = ne;
      else if (ne != numel[t]) { err = "Size mismatch"; break; }
      if (x.scalar_type() != list_dtype) { err = "tensors of one list must share a dtype"; break; }
      ptrs[d * T + t] = reinterpret_cast<uint64_t>(x.data_ptr());
    }
    Py_DECREF(inner);
    if (not_tensor) { Py_DECREF(outer); Py_RETURN_NONE; }
    if (err) { Py_DECREF(outer); return fail(err); }
  }
  Py_DECREF(outer);
  for (Py_ssize_t d = 0; d < depth; d++)
    if (codes[d] < 0) return fail("multi_tensor_apply not implemented for this dtype");
  PyObject* ct = PyTuple_New(depth);
  for (Py_ssize_t d = 0; d < depth; d++) PyTuple_SET_ITEM(ct, d, PyLong_FromLong(codes[d]));
  return Py_BuildValue("(nnNi)", T, depth, ct, (int)ref_device.index());
}

// check_rows(rows, name0, name1, grad_ptrs_addr, dtype_code) -> 0 | 1 | 2
//   rows: list of (param, state_dict, state0, state1) as cached by apex.optimizers._common.GroupCache.
//   0: every param has a dense, contiguous gradient of the cached dtype at the cached address AND state_dict[name0/1] are
//      still the cached tensors  -> the prepared [g, p, s0, s1] handle may be used as it is;
//   1: the state is unchanged but a gradient is missing, moved or changed type -> re-read the gradients;
//   2: a state tensor was replaced -> the cache entry is stale.
// One C loop instead of a Python loop with two dict look-ups and an attribute read per parameter.
PyObject* check_rows(PyObject*, PyObject* args) {
  PyObject *rows, *n0, *n1;
  unsigned long long addr;
  int code;
  if (!PyArg_ParseTuple(args, "OOOKi", &rows, &n0, &n1, &addr, &code)) return nullptr;
  if (!PyList_Check(rows)) return fail("check_rows: rows must be a list");
  const uint64_t* want = reinterpret_cast<const uint64_t*>(addr);
  const Py_ssize_t n = PyList_GET_SIZE(rows);
  long rc = 0;
  for (Py_ssize_t i = 0; i < n; i++) {
    PyObject* row = PyList_GET_ITEM(rows, i);
    if (!PyTuple_Check(row) || PyTuple_GET_SIZE(row) != 4) return fail("check_rows: malformed row");
    PyObject* p = PyTuple_GET_ITEM(row, 0);
    PyObject* st = PyTuple_GET_ITEM(row, 1);
    if (!THPVariable_Check(p) || !PyDict_Check(st)) return fail("check_rows: malformed row");
    if (PyDict_GetItem(st, n0) != PyTuple_GET_ITEM(row, 2) || PyDict_GetItem(st, n1) != PyTuple_GET_ITEM(row, 3))
      return PyLong_FromLong(2);
    if (rc) continue;
    const at::Tensor& x = THPVariable_Unpack(p);
    const at::Tensor& g = x.grad();
    if (!g.defined() || g.is_sparse() || dtype_code(g.scalar_type()) != code || !g.is_contiguous() ||
        reinterpret_cast<uint64_t>(g.data_ptr()) != want[i])
      rc = 1;
  }
  return PyLong_FromLong(rc);
}

PyMethodDef methods[] = {
    {"check_rows", check_rows, METH_VARARGS, "are the cached gradient addresses and state tensors of a parameter group still valid?"},
    {"pack_lists", pack_lists, METH_VARARGS, "tensor lists -> raw pointer / numel arrays"},
    {nullptr, nullptr, 0, nullptr}};

PyModuleDef module = {PyModuleDef_HEAD_INIT, "_b200mt_pack", "fast tensor-list packer for libb200mt", -1, methods};

}  // namespace

PyMODINIT_FUNC PyInit__b200mt_pack(void) { return PyModule_Create(&module); }
