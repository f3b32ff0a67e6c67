// Host-side writers/readers of the reduced operators in the reference's on-disk formats, so
// that `./ITHACAoutput/Matrices/*` written from the C ABI's output buffers is read back by an unmodified ITHACA-FV
// (steadyNS.C:384-467 ReadDenseMatrix / ReadDenseTensor cache branch) and by the reference's python tooling.
// No device code here; compiled by nvcc only because the whole library is.
//   SaveDenseMatrix / ReadDenseMatrix   src/ITHACA_CORE/ITHACAstream/ITHACAstream.H:722-762
//   SaveDenseTensor / ReadDenseTensor   ITHACAstream.H:765-809
//   exportMatrix (python/matlab/eigen)  src/ITHACA_CORE/ITHACAstream/ITHACAstream.C:54-147
//   exportTensor                        ITHACAstream.C:263-358
//   cnpy::save (matrix, rank-3 tensor)  src/ITHACA_CORE/ITHACAstream/cnpy.C:524-568, header cnpy.H:454-488
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <string>
#include <sys/stat.h>
#include <vector>

namespace ib200 {
void set_error(ithaca_b200* h, const std::string& msg);
}
using ib200::set_error;

static int io_fail(const char* what, const char* path)
{
    set_error(nullptr, std::string(what) + " '" + (path ? path : "(null)") + "': " + strerror(errno));
    return IB200_ERR_ARG;
}

// Foam::mkDir semantics: create the folder and its parents, succeed when it is already there.
static int mkdir_p(const std::string& folder)
{
    if (folder.empty()) return 0;
    std::string cur;
    for (size_t i = 0; i <= folder.size(); i++) {
        if (i == folder.size() || folder[i] == '/') {
            if (!cur.empty() && cur != "." && cur != "..") {
                if (mkdir(cur.c_str(), 0777) != 0 && errno != EEXIST) return -1;
            }
        }
        if (i < folder.size()) cur += folder[i];
    }
    return 0;
}

static std::string join(const char* folder, const std::string& file)
{
    std::string f = folder ? folder : "";
    return f + "/" + file;
}

extern "C" {

int ithaca_b200_save_dense_matrix(const char* path, const double* A, int64_t rows, int64_t cols)
{
    if (!path || rows < 0 || cols < 0 || (!A && rows * cols > 0)) {
        set_error(nullptr, "save_dense_matrix: bad argument");
        return IB200_ERR_ARG;
    }
    FILE* f = fopen(path, "wb");
    if (!f) return io_fail("cannot open", path);
    // Eigen::Index is std::ptrdiff_t: two 8-byte integers, then the column-major data
    bool ok = fwrite(&rows, sizeof(int64_t), 1, f) == 1 && fwrite(&cols, sizeof(int64_t), 1, f) == 1;
    const size_t n = (size_t)rows * (size_t)cols;
    ok = ok && (n == 0 || fwrite(A, sizeof(double), n, f) == n);
    if (fclose(f) != 0) ok = false;
    return ok ? IB200_OK : io_fail("short write to", path);
}

int ithaca_b200_read_dense_matrix(const char* path, int64_t* rows, int64_t* cols, double* A, int64_t capacity)
{
    if (!path || !rows || !cols) {
        set_error(nullptr, "read_dense_matrix: bad argument");
        return IB200_ERR_ARG;
    }
    FILE* f = fopen(path, "rb");
    if (!f) return io_fail("file does not exist, try to rerun the Offline Stage:", path);
    int64_t r = 0, c = 0;
    bool ok = fread(&r, sizeof(int64_t), 1, f) == 1 && fread(&c, sizeof(int64_t), 1, f) == 1 && r >= 0 && c >= 0;
    if (ok) {
        *rows = r;
        *cols = c;
        if (A) {
            const size_t n = (size_t)r * (size_t)c;
            if ((int64_t)n > capacity) {
                fclose(f);
                set_error(nullptr, "read_dense_matrix: buffer too small");
                return IB200_ERR_ARG;
            }
            ok = n == 0 || fread(A, sizeof(double), n, f) == n;
        }
    }
    fclose(f);
    if (!ok) {
        set_error(nullptr, std::string("read_dense_matrix: truncated file '") + path + "'");
        return IB200_ERR_ARG;
    }
    return IB200_OK;
}

int ithaca_b200_save_dense_tensor(const char* path, const double* T, int64_t d0, int64_t d1, int64_t d2)
{
    if (!path || d0 < 0 || d1 < 0 || d2 < 0 || (!T && d0 * d1 * d2 > 0)) {
        set_error(nullptr, "save_dense_tensor: bad argument");
        return IB200_ERR_ARG;
    }
    FILE* f = fopen(path, "wb");
    if (!f) return io_fail("cannot open", path);
    // Eigen::Tensor<double,3>::Dimensions = DSizes<Eigen::Index,3>: three 8-byte integers, then the data, first index fastest
    const int64_t dim[3] = {d0, d1, d2};
    const size_t n = (size_t)d0 * (size_t)d1 * (size_t)d2;
    bool ok = fwrite(dim, sizeof(int64_t), 3, f) == 3 && (n == 0 || fwrite(T, sizeof(double), n, f) == n);
    if (fclose(f) != 0) ok = false;
    return ok ? IB200_OK : io_fail("short write to", path);
}

int ithaca_b200_read_dense_tensor(const char* path, int64_t* dims3, double* T, int64_t capacity)
{
    if (!path || !dims3) {
        set_error(nullptr, "read_dense_tensor: bad argument");
        return IB200_ERR_ARG;
    }
    FILE* f = fopen(path, "rb");
    if (!f) return io_fail("cannot open", path);
    int64_t dim[3] = {0, 0, 0};
    bool ok = fread(dim, sizeof(int64_t), 3, f) == 3 && dim[0] >= 0 && dim[1] >= 0 && dim[2] >= 0;
    if (ok) {
        memcpy(dims3, dim, sizeof(dim));
        if (T) {
            const size_t n = (size_t)dim[0] * (size_t)dim[1] * (size_t)dim[2];
            if ((int64_t)n > capacity) {
                fclose(f);
                set_error(nullptr, "read_dense_tensor: buffer too small");
                return IB200_ERR_ARG;
            }
            ok = n == 0 || fread(T, sizeof(double), n, f) == n;
        }
    }
    fclose(f);
    if (!ok) {
        set_error(nullptr, std::string("read_dense_tensor: truncated file '") + path + "'");
        return IB200_ERR_ARG;
    }
    return IB200_OK;
}

int ithaca_b200_save_npy(const char* path, const double* data, int ndim, const int64_t* dims)
{
    if (!path || !dims || ndim < 1 || ndim > 3) {
        set_error(nullptr, "save_npy: ndim must be 1, 2 or 3");
        return IB200_ERR_ARG;
    }
    size_t n = 1;
    for (int k = 0; k < ndim; k++) {
        if (dims[k] < 0) {
            set_error(nullptr, "save_npy: negative dimension");
            return IB200_ERR_ARG;
        }
        n *= (size_t)dims[k];
    }
    // header exactly as cnpy writes it: version 1.0, little-endian '<f8', C order, padded so that the 10-byte
    // preamble + dictionary is a multiple of 16 (a full extra 16 when it already is), newline-terminated
    std::string dict = "{'descr': '<f8', 'fortran_order': False, 'shape': (";
    dict += std::to_string((unsigned long long)dims[0]);
    for (int k = 1; k < ndim; k++) dict += ", " + std::to_string((unsigned long long)dims[k]);
    if (ndim == 1) dict += ",";
    dict += "), }";
    const int remainder = 16 - (int)((10 + dict.size()) % 16);
    dict.append((size_t)remainder, ' ');
    dict.back() = '\n';
    // cnpy::save copies the Eigen object (first index fastest) into row-major order (last index fastest)
    std::vector<double> rowmajor(n);
    const int64_t d0 = dims[0], d1 = ndim > 1 ? dims[1] : 1, d2 = ndim > 2 ? dims[2] : 1;
    for (int64_t i = 0; i < d0; i++)
        for (int64_t j = 0; j < d1; j++)
            for (int64_t k = 0; k < d2; k++) rowmajor[(size_t)((i * d1 + j) * d2 + k)] = data[(size_t)(i + d0 * (j + d1 * k))];
    FILE* f = fopen(path, "wb");
    if (!f) return io_fail("cannot open", path);
    const unsigned char pre[8] = {0x93, 'N', 'U', 'M', 'P', 'Y', 0x01, 0x00};
    const uint16_t hl = (uint16_t)dict.size();
    bool ok = fwrite(pre, 1, 8, f) == 8 && fwrite(&hl, 2, 1, f) == 1 && fwrite(dict.data(), 1, dict.size(), f) == dict.size() &&
              (n == 0 || fwrite(rowmajor.data(), sizeof(double), n, f) == n);
    if (fclose(f) != 0) ok = false;
    return ok ? IB200_OK : io_fail("short write to", path);
}

// One matrix in the text format `type` of ITHACAstream::exportMatrix: <folder>/<name>_mat.{py,m,txt}.
static int export_matrix_impl(const char* folder, const std::string& name, int type, const double* A, int64_t rows,
                              int64_t cols)
{
    if (mkdir_p(folder ? folder : "") != 0) return io_fail("cannot create", folder);
    const char* est = type == IB200_EXPORT_PYTHON ? ".py" : type == IB200_EXPORT_MATLAB ? ".m" : ".txt";
    const std::string path = join(folder, name + "_mat" + est);
    FILE* f = fopen(path.c_str(), "w");
    if (!f) return io_fail("cannot open", path.c_str());
    if (type == IB200_EXPORT_PYTHON) {
        fprintf(f, "%s=np.array([", name.c_str());
        for (int64_t i = 0; i < rows; i++) {
            for (int64_t j = 0; j < cols; j++) fprintf(f, j == 0 ? "[%.10g" : ",%.10g", A[i + rows * j]);
            if (i != rows - 1) fprintf(f, "],\n");
        }
        fprintf(f, "]])\n");
    } else if (type == IB200_EXPORT_MATLAB) {
        fprintf(f, "%s=[", name.c_str());
        for (int64_t i = 0; i < rows; i++) {
            for (int64_t j = 0; j < cols; j++) fprintf(f, " %.10g", A[i + rows * j]);
            if (i != rows - 1) fprintf(f, ";\n");
        }
        fprintf(f, "];\n");
    } else {
        for (int64_t i = 0; i < rows; i++) {
            for (int64_t j = 0; j < cols; j++) fprintf(f, j == 0 ? "%.20g" : " %.20g", A[i + rows * j]);
            if (i != rows - 1) fprintf(f, "\n");
        }
    }
    if (fclose(f) != 0) return io_fail("short write to", path.c_str());
    return IB200_OK;
}

int ithaca_b200_export_matrix(const char* folder, const char* name, int type, const double* A, int64_t rows, int64_t cols)
{
    if (!name || rows < 0 || cols < 0 || (!A && rows * cols > 0) || type < IB200_EXPORT_PYTHON || type > IB200_EXPORT_EIGEN) {
        // the reference asserts on the extension (ITHACAstream.C:59-63)
        set_error(nullptr, "The extension was not implemented. Check the list of possible extensions.");
        return IB200_ERR_ARG;
    }
    return export_matrix_impl(folder, name, type, A, rows, cols);
}

int ithaca_b200_export_tensor(const char* folder, const char* name, int type, const double* T, int64_t d0, int64_t d1,
                              int64_t d2)
{
    if (!name || d0 < 0 || d1 < 0 || d2 < 0 || (!T && d0 * d1 * d2 > 0) || type < IB200_EXPORT_PYTHON ||
        type > IB200_EXPORT_EIGEN) {
        set_error(nullptr, "The extension was not implemented. Check the list of possible extensions.");
        return IB200_ERR_ARG;
    }
    // slice i = Eigen::SliceFromTensor(tensor, 0, i): the d1 x d2 matrix T(i, :, :)
    auto at = [&](int64_t i, int64_t j, int64_t k) { return T[(size_t)(i + d0 * (j + d1 * k))]; };
    if (type == IB200_EXPORT_EIGEN) {
        std::vector<double> slice((size_t)d1 * (size_t)d2);
        for (int64_t i = 0; i < d0; i++) {
            for (int64_t k = 0; k < d2; k++)
                for (int64_t j = 0; j < d1; j++) slice[(size_t)(j + d1 * k)] = at(i, j, k);
            int rc = export_matrix_impl(folder, std::string(name) + std::to_string((long long)i), type, slice.data(), d1, d2);
            if (rc) return rc;
        }
        return IB200_OK;
    }
    if (mkdir_p(folder ? folder : "") != 0) return io_fail("cannot create", folder);
    const std::string path = join(folder, std::string(name) + "_mat" + (type == IB200_EXPORT_PYTHON ? ".py" : ".m"));
    FILE* f = fopen(path.c_str(), "w");
    if (!f) return io_fail("cannot open", path.c_str());
    if (type == IB200_EXPORT_PYTHON) {
        fprintf(f, "%s=np.zeros([%lld,%lld,%lld])\n", name, (long long)d0, (long long)d1, (long long)d2);
        for (int64_t i = 0; i < d0; i++) {
            fprintf(f, "%s[%lld,:,:]=np.array([", name, (long long)i);
            for (int64_t j = 0; j < d1; j++) {
                for (int64_t k = 0; k < d2; k++) fprintf(f, k == 0 ? "[%.10g" : ",%.10g", at(i, j, k));
                if (j != d1 - 1) fprintf(f, "],\n");
            }
            fprintf(f, "]])\n\n");
        }
    } else {
        for (int64_t i = 0; i < d0; i++) {
            fprintf(f, "%s(%lld,:,:)=[", name, (long long)(i + 1));
            for (int64_t j = 0; j < d1; j++) {
                for (int64_t k = 0; k < d2; k++) fprintf(f, " %.10g", at(i, j, k));
                if (j != d1 - 1) fprintf(f, ";\n");
            }
            fprintf(f, "];\n");
        }
    }
    if (fclose(f) != 0) return io_fail("short write to", path.c_str());
    return IB200_OK;
}

}  // extern "C"
