// Host side of the plugin boundary, native: one C++ walk over the reference's List[List[Tensor]]
// (base_defense.py:31-35) that validates it and fills the engine's pointer table.
//
// Why: Updates.__init__ / _ptr_tables touch every tensor handle from Python -- numel, is_cuda, dtype,
// is_contiguous, requires_grad, data_ptr: ~0.8 us per tensor.  That is 0.4 ms at n = 500 clients with one tensor
// each, and 13 ms for 100 clients x the 161 parameter tensors of the reference's ResNet-50 (cub200_cnn.py:26-29):
// several times the 2 ms the median kernel takes.  The same walk here costs ~20 ns per tensor.
//
// walk(client_updates, table_address) -> (status, n, P, on_host, device_index)
//   status 0: every tensor is fp32, contiguous, detached or grad-free, on one device type, shapes consistent;
//             table_address (a uint64 array of n*P entries owned by the caller) holds the data pointers.
//   status 1: something needs the general path (a dtype / layout conversion, requires_grad, or an error the
//             Python constructor reports with its own message); the table is not valid.
// Built by csrc/Makefile into flagg/_hostwalk*.so; flagg.engine uses it when present and falls back to the
// Python walk (same checks, same errors) when it is not -- this is host bookkeeping, not a compute path.
#include <Python.h>
#include <torch/csrc/autograd/python_variable.h>
#include <ATen/core/Tensor.h>

static PyObject* walk(PyObject*, PyObject* args) {
    PyObject* updates;
    unsigned long long table_addr;
    if (!PyArg_ParseTuple(args, "OK", &updates, &table_addr)) return nullptr;
    uint64_t* table = reinterpret_cast<uint64_t*>(static_cast<uintptr_t>(table_addr));
    if (!PyList_Check(updates) && !PyTuple_Check(updates)) return Py_BuildValue("(iiiii)", 1, 0, 0, 0, -1);
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(updates);
    if (n == 0) return Py_BuildValue("(iiiii)", 1, 0, 0, 0, -1);
    PyObject** rows = PySequence_Fast_ITEMS(updates);
    Py_ssize_t P = -1;
    bool on_host = false;
    int dev_index = -1;
    std::vector<int64_t> numel;
    for (Py_ssize_t i = 0; i < n; ++i) {
        PyObject* row = rows[i];
        if (!PyList_Check(row) && !PyTuple_Check(row)) return Py_BuildValue("(iiiii)", 1, 0, 0, 0, -1);
        const Py_ssize_t p_i = PySequence_Fast_GET_SIZE(row);
        if (i == 0) {
            P = p_i;
            if (P == 0) return Py_BuildValue("(iiiii)", 1, 0, 0, 0, -1);
            numel.resize(P);
        } else if (p_i != P) {
            return Py_BuildValue("(iiiii)", 1, 0, 0, 0, -1);
        }
        PyObject** ts = PySequence_Fast_ITEMS(row);
        for (Py_ssize_t p = 0; p < P; ++p) {
            if (!THPVariable_Check(ts[p])) return Py_BuildValue("(iiiii)", 1, 0, 0, 0, -1);
            const at::Tensor& t = THPVariable_Unpack(ts[p]);
            if (!t.defined() || t.scalar_type() != at::kFloat || !t.is_contiguous() || t.requires_grad() ||
                t.layout() != at::kStrided || t.is_conj() || t.is_neg())
                return Py_BuildValue("(iiiii)", 1, 0, 0, 0, -1);
            const bool host = !t.is_cuda();
            if (host && !t.is_cpu()) return Py_BuildValue("(iiiii)", 1, 0, 0, 0, -1);
            if (i == 0 && p == 0) {
                on_host = host;
                dev_index = host ? -1 : int(t.get_device());
            } else if (host != on_host || (!host && int(t.get_device()) != dev_index)) {
                return Py_BuildValue("(iiiii)", 1, 0, 0, 0, -1);
            }
            if (i == 0) numel[p] = t.numel();
            else if (t.numel() != numel[p]) return Py_BuildValue("(iiiii)", 1, 0, 0, 0, -1);
            table[size_t(i) * size_t(P) + size_t(p)] = reinterpret_cast<uint64_t>(t.data_ptr());
        }
    }
    return Py_BuildValue("(iiiii)", 0, int(n), int(P), on_host ? 1 : 0, dev_index);
}

static PyMethodDef methods[] = {{"walk", walk, METH_VARARGS, "validate List[List[Tensor]] and fill the pointer table"},
                                {nullptr, nullptr, 0, nullptr}};
static struct PyModuleDef moddef = {PyModuleDef_HEAD_INIT, "_hostwalk", "native walk over the plugin API's tensor lists", -1, methods};
PyMODINIT_FUNC PyInit__hostwalk(void) { return PyModule_Create(&moddef); }
