// Host C++ side of the rows after the reward path: the slice writer with its
// metadata.bin sidecar and the decoder, over the C ABI of libicr_b200.so.  Replaces, as drop-ins:
//   write_slices                                     src/image_slicer.cpp:81-130
//   write_metadata_binary / read_metadata_binary     src/metadata.cpp:4-80
//   reassemble                                       src/reassemble.cpp:30-120
// The device cuts every slice of the batch in one pass and pastes on the decoder side; this file moves bytes, encodes
// and decodes the PNG container (zlib, host threads) and writes files.
#include <ATen/ops/_unique.h>
#include <c10/cuda/CUDAGuard.h>

#include <algorithm>
#include <atomic>
#include <fstream>
#include <thread>

#include "host_internal.hpp"
#include "include/image_slicer.h"
#include "include/reassemble.h"
#include "png_io.hpp"

using icrh::current_stream;
using icrh::guarded;
using icrh::label_code;
using icrh::throw_runtime;
using icrh::WsHolder;

namespace {

constexpr int PNG_LEVEL = 4;            // COMPRESSION_LEVEL, include/configuration.h:11

void check_rc(int rc) {
    if (rc == ICR_ECUDA) throw_runtime();
    TORCH_CHECK(rc == ICR_OK, icr_last_error());
}

}  // namespace

// ----------------------------------------------------------------------------------------------------- metadata.bin
namespace {
void put_le(std::vector<char>& out, uint64_t v, int bytes) {
    for (int i = 0; i < bytes; ++i) out.push_back((char)((v >> (8 * i)) & 0xffu));
}
uint64_t get_le(const unsigned char* p, int bytes) {
    uint64_t v = 0;
    for (int i = 0; i < bytes; ++i) v |= (uint64_t)p[i] << (8 * i);
    return v;
}
}  // namespace

void write_metadata_binary(const std::vector<SliceMetadata>& metadata, const std::string& path, uint32_t image_width,
                           uint32_t image_height) {
    std::vector<char> out;
    out.reserve(icr_sidecar::kHeaderBytes + metadata.size() * (icr_sidecar::kRecordBytes + 16));
    put_le(out, icr_sidecar::kMagic, 4);
    put_le(out, (uint32_t)metadata.size(), 4);
    put_le(out, image_width, 4);
    put_le(out, image_height, 4);
    for (const SliceMetadata& m : metadata) {
        const int32_t fields[5] = {m.label, m.x, m.y, m.width, m.height};
        for (int32_t f : fields) put_le(out, (uint32_t)f, 4);
        put_le(out, (uint16_t)m.filename.size(), 2);                  // the reference truncates the length to 16 bits too
        out.insert(out.end(), m.filename.begin(), m.filename.end());
    }
    std::ofstream f(path, std::ios::binary | std::ios::trunc);
    f.write(out.data(), (std::streamsize)out.size());
}

std::vector<SliceMetadata> read_metadata_binary(const std::string& path, uint32_t& image_width, uint32_t& image_height) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("Cannot open metadata file");
    unsigned char head[icr_sidecar::kHeaderBytes] = {};
    f.read((char*)head, sizeof head);
    if (get_le(head, 4) != icr_sidecar::kMagic) throw std::runtime_error("Invalid metadata file (magic mismatch)");
    const uint32_t count = (uint32_t)get_le(head + 4, 4);
    image_width = (uint32_t)get_le(head + 8, 4);
    image_height = (uint32_t)get_le(head + 12, 4);
    std::vector<SliceMetadata> records;
    for (uint32_t i = 0; i < count && f; ++i) {
        unsigned char fixed[icr_sidecar::kRecordBytes] = {};
        f.read((char*)fixed, sizeof fixed);
        SliceMetadata m;
        m.label = (int32_t)get_le(fixed, 4);
        m.x = (int32_t)get_le(fixed + 4, 4);
        m.y = (int32_t)get_le(fixed + 8, 4);
        m.width = (int32_t)get_le(fixed + 12, 4);
        m.height = (int32_t)get_le(fixed + 16, 4);
        m.filename.resize((size_t)get_le(fixed + 20, 2));
        if (!m.filename.empty()) f.read(&m.filename[0], (std::streamsize)m.filename.size());
        records.push_back(std::move(m));
    }
    return records;
}

// ------------------------------------------------------------------------------------------------------ image_slicer
SliceBatch slice_images_batched(const torch::Tensor& images_bchw_f32, const torch::Tensor& labels, bool swap_rb) {
    TORCH_CHECK(images_bchw_f32.is_cuda() && labels.is_cuda(), "images and labels must be CUDA");
    TORCH_CHECK(images_bchw_f32.dim() == 4 && images_bchw_f32.size(1) == 3, "images must be [B,3,H,W]");
    TORCH_CHECK(images_bchw_f32.scalar_type() == torch::kFloat32, "images must be float32");
    TORCH_CHECK(labels.dim() == 3 || labels.dim() == 4, "labels must be [B,H,W] or [B,S,H,W]");
    TORCH_CHECK(c10::isIntegralType(labels.scalar_type(), false), "labels must be an integer tensor");
    torch::NoGradGuard ng;
    c10::cuda::CUDAGuard dg(images_bchw_f32.device());
    const int64_t B = images_bchw_f32.size(0), H = images_bchw_f32.size(2), W = images_bchw_f32.size(3);
    const int64_t S = labels.dim() == 4 ? labels.size(1) : 1;
    TORCH_CHECK(labels.size(0) == B && labels.size(-2) == H && labels.size(-1) == W, "labels shape mismatch");
    torch::Tensor img = images_bchw_f32.contiguous();
    torch::Tensor labs = labels;
    if (labs.scalar_type() != torch::kInt64 && labs.scalar_type() != torch::kInt32) labs = labs.to(torch::kInt64);
    labs = labs.contiguous();
    const int64_t V = B * S;
    const icr_stream st = current_stream();
    int Kmax = (int)std::min<int64_t>(H * W, icrh::KMAX_FIRST);
    bool rank = false, ranked = false;
    torch::Tensor raw_labs = labs;                               // keys are reported in the caller's label values
    WsHolder* hp = nullptr;
    for (int attempt = 0;; ++attempt) {                          // grown tables / ranked labels, as compute_rewards_batched does
        if (rank && !ranked) { labs = icrh::rank_labels(labs.view({V, H, W})).view(labs.sizes()); ranked = true; }
        hp = &icrh::workspace_for(img.device().index(), (int)V, (int)H, (int)W, Kmax);
        check_rc(icr_slices_prepare(img.data_ptr<float>(), labs.data_ptr(), label_code(labs), (int)B, (int)S, (int)H, (int)W, hp->ws, st));
        int bits = 0, maxk = 0;
        const int rc = icr_check_detail(hp->ws, st, nullptr, &bits, &maxk);
        if (rc == ICR_EDATA && attempt < 3 && icrh::plan_retry(bits, maxk, H * W, Kmax, rank)) continue;
        check_rc(rc);
        break;
    }
    WsHolder& h = *hp;
    const int32_t *pK, *pb; const long long* pk; int km = 0;
    icr_workspace_tables(h.ws, &pK, nullptr, &pb, nullptr, &pk, &km);
    auto i32 = torch::TensorOptions().device(img.device()).dtype(torch::kInt32);
    auto i64 = i32.dtype(torch::kInt64);
    SliceBatch out;
    out.K = torch::from_blob((void*)pK, {V}, i32).clone();
    const int64_t kcap = std::max<int64_t>(out.K.max().item<int>(), 1);          // rows in use: the tables are H*W rows wide
    out.bboxes = torch::from_blob((void*)pb, {V, (int64_t)km, 4}, i32).slice(1, 0, kcap).clone();
    out.keys = torch::from_blob((void*)pk, {V, (int64_t)km}, i64).slice(1, 0, kcap).clone();
    if (ranked) {        // the tables hold ranks: report the caller's label values, ascending (what the ranks index)
        const torch::Tensor raw = raw_labs.view({V, H * W});
        for (int64_t v = 0; v < V; ++v) {
            const torch::Tensor uniq = std::get<0>(at::_unique(raw[v], /*sorted=*/true, /*return_inverse=*/false)).to(torch::kInt64);
            out.keys[v].slice(0, 0, uniq.numel()).copy_(uniq);
        }
    }
    const int64_t nseg_cap = out.K.sum().item<int64_t>() + 1;
    torch::Tensor offsets = torch::zeros({nseg_cap}, i64);
    torch::Tensor summary = torch::zeros({2}, i64);
    check_rc(icr_slices_plan(h.ws, (unsigned long long*)offsets.data_ptr<int64_t>(), nseg_cap,
                             (unsigned long long*)summary.data_ptr<int64_t>(), st));
    const torch::Tensor sh = summary.cpu();                                       // the one 16-byte read-back that sizes the buffer
    const int64_t nseg = sh[0].item<int64_t>(), nbytes = sh[1].item<int64_t>();
    TORCH_CHECK(nseg + 1 == nseg_cap, "slice plan disagrees with the segment counts");
    out.bytes = torch::empty({std::max<int64_t>(nbytes, 4)}, i32.dtype(torch::kUInt8));
    check_rc(icr_slices_emit(h.ws, (const unsigned long long*)offsets.data_ptr<int64_t>(), out.bytes.data_ptr<uint8_t>(), (size_t)nbytes,
                             swap_rb ? ICR_ORDER_SWAPPED : ICR_ORDER_PLANES, st));
    out.bytes = out.bytes.slice(0, 0, nbytes);
    out.offsets = offsets;
    return out;
}

torch::Tensor slice_image(const torch::Tensor& input_chw_f32, const torch::Tensor& mask, int64_t label, SliceBox& out_box) {
    TORCH_CHECK(input_chw_f32.dim() == 3 || (input_chw_f32.dim() == 4 && input_chw_f32.size(0) == 1), "input must be [3,H,W] or [1,3,H,W]");
    TORCH_CHECK(mask.dim() == 2, "mask must be [H,W]");
    const torch::Tensor img = input_chw_f32.dim() == 3 ? input_chw_f32.unsqueeze(0) : input_chw_f32;
    // the label pass of the whole map (Rules L, S) and its slices are one device pass; the one asked for is picked out
    const SliceBatch sb = slice_images_batched(img, mask.unsqueeze(0), false);
    const int K = sb.K.cpu()[0].item<int>();
    const torch::Tensor keys = sb.keys[0].slice(0, 0, K).cpu();
    const int64_t* pk = keys.data_ptr<int64_t>();
    const int64_t* hit = std::lower_bound(pk, pk + K, label);           // keys ascend (Rule L)
    if (hit == pk + K || *hit != label) return torch::Tensor();        // :55-57
    const int64_t k = hit - pk;
    const torch::Tensor bb = sb.bboxes[0][k].cpu(), off = sb.offsets.slice(0, k, k + 2).cpu();
    const int x0 = bb[0].item<int>(), y0 = bb[1].item<int>(), w = bb[2].item<int>() - x0 + 1, h = bb[3].item<int>() - y0 + 1;
    out_box.x = x0; out_box.y = y0; out_box.width = w; out_box.height = h;     // :68-71
    return sb.bytes.slice(0, off[0].item<int64_t>(), off[1].item<int64_t>()).view({h, w, 4}).clone();
}

bool write_slices(const torch::Tensor& input_chw_f32, const torch::Tensor& mask, const std::filesystem::path& output_path,
                  const std::filesystem::path& file_directory_name) {
    TORCH_CHECK(input_chw_f32.dim() == 3 || (input_chw_f32.dim() == 4 && input_chw_f32.size(0) == 1), "input must be [3,H,W] or [1,3,H,W]");
    TORCH_CHECK(mask.dim() == 2, "mask must be [H,W]");
    const auto dir = output_path / file_directory_name;
    if (!std::filesystem::exists(dir)) std::filesystem::create_directories(dir);     // :87-89
    const torch::Tensor img = input_chw_f32.dim() == 3 ? input_chw_f32.unsqueeze(0) : input_chw_f32;
    const SliceBatch sb = slice_images_batched(img, mask.unsqueeze(0), false);
    const int64_t H = img.size(2), W = img.size(3);
    // one copy of everything the host needs
    const torch::Tensor bytes = sb.bytes.cpu(), offsets = sb.offsets.cpu(), bboxes = sb.bboxes.cpu(), keys = sb.keys.cpu();
    const int K = sb.K.cpu()[0].item<int>();
    const uint8_t* pb = bytes.data_ptr<uint8_t>();
    const int64_t* po = offsets.data_ptr<int64_t>();
    const int32_t* pbb = bboxes.data_ptr<int32_t>();
    const int64_t* pk = keys.data_ptr<int64_t>();
    std::vector<SliceMetadata> metadata(K);
    std::atomic<int> next{0};
    std::atomic<bool> success{true};
    const unsigned nthreads = std::max(1u, std::min<unsigned>(std::thread::hardware_concurrency(), (unsigned)std::max(K, 1)));   // :93
    auto work = [&] {
        for (int k = next.fetch_add(1); k < K; k = next.fetch_add(1)) {
            const int x0 = pbb[4 * k], y0 = pbb[4 * k + 1], w = pbb[4 * k + 2] - x0 + 1, h = pbb[4 * k + 3] - y0 + 1;
            SliceMetadata& m = metadata[k];
            m.label = (int)pk[k];                                                   // :112
            m.filename = "slice_" + std::to_string(pk[k]) + ".png";                 // :107 (IMAGE_FORMAT = "png")
            m.x = x0; m.y = y0; m.width = w; m.height = h;
            const std::vector<uint8_t> png = icr_png::encode_rgba(pb + po[k], w, h, PNG_LEVEL);
            if (!icr_png::write_file((dir / m.filename).string(), png)) success = false;    // :108
        }
    };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < nthreads; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    write_metadata_binary(metadata, (dir / "metadata.bin").string(), (uint32_t)W, (uint32_t)H);      // :127
    return success;
}

// -------------------------------------------------------------------------------------------------------- reassemble
torch::Tensor reassemble_to_tensor(const std::filesystem::path& image_path, int device) {
    uint32_t w = 0, h = 0;
    const std::vector<SliceMetadata> meta = read_metadata_binary((image_path / "metadata.bin").string(), w, h);
    if (meta.empty()) throw std::runtime_error("No slices in metadata");             // :46-49
    c10::cuda::CUDAGuard dg((c10::DeviceIndex)device);
    auto u8 = torch::TensorOptions().device(torch::Device(torch::kCUDA, device)).dtype(torch::kUInt8);
    torch::Tensor canvas = torch::zeros({(int64_t)h, (int64_t)w, 4}, u8);            // :50
    const icr_stream st = current_stream();
    std::vector<torch::Tensor> keep;                                                 // device copies live until the stream is done
    for (const SliceMetadata& m : meta) {
        if (m.filename.empty()) continue;                                            // :53-56
        std::vector<uint8_t> file, rgba;
        if (!icr_png::read_file((image_path / m.filename).string(), file) && !icr_png::read_file(m.filename, file)) continue;   // :58-62
        int sw = 0, sh = 0;
        std::string err;
        if (!icr_png::decode_rgba(file, rgba, sw, sh, err)) continue;
        if (m.x < 0 || m.y < 0) continue;
        // the actual slice size is used, as :66-71 warns
        torch::Tensor d = torch::from_blob(rgba.data(), {(int64_t)sh, (int64_t)sw, 4}, torch::kUInt8).to(canvas.device());
        keep.push_back(d);
        check_rc(icr_paste_slice(canvas.data_ptr<uint8_t>(), (int)w, (int)h, d.data_ptr<uint8_t>(), m.x, m.y, sw, sh, st));
    }
    c10::cuda::getCurrentCUDAStream().synchronize();
    return canvas;
}

bool reassemble(std::filesystem::path image_path, std::filesystem::path out_filename) {
    torch::Tensor canvas;
    try {
        canvas = reassemble_to_tensor(image_path, 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "Error reading metadata: %s\n", e.what());                   // :38-41
        return false;
    }
    const torch::Tensor c = canvas.cpu().contiguous();
    const std::vector<uint8_t> png = icr_png::encode_rgba(c.data_ptr<uint8_t>(), (int)c.size(1), (int)c.size(0), PNG_LEVEL);
    return icr_png::write_file(out_filename.string(), png);                          // :112-115
}

// ---------------------------------------------------------------------------------------------------- test hooks
extern "C" {

int icrh_write_slices(const float* image_dev, const void* mask_dev, int mask_i64, int H, int W, int device, const char* output_path,
                      const char* dir_name) {
    int ok = 0;
    const int rc = guarded([&] {
        auto dev = torch::Device(torch::kCUDA, device);
        auto img = torch::from_blob((void*)image_dev, {3, H, W}, torch::TensorOptions().dtype(torch::kFloat32).device(dev));
        auto mask = torch::from_blob((void*)mask_dev, {H, W}, torch::TensorOptions().dtype(mask_i64 ? torch::kInt64 : torch::kInt32).device(dev));
        ok = write_slices(img, mask, output_path, dir_name) ? 1 : 0;
    });
    return rc ? rc : (ok ? 0 : 4);
}

// slice_image for one label value: box_out = {x, y, w, h}; the slice goes to slice_out_dev (capacity bytes).  Returns 0,
// or -1 when the label does not occur (empty result, as the reference's empty cv::Mat).
int icrh_slice_image(const float* image_dev, const void* mask_dev, int mask_i64, int H, int W, int device, long long label,
                     int32_t* box_out, uint8_t* slice_out_dev, long long capacity) {
    int found = 0;
    const int rc = guarded([&] {
        c10::cuda::CUDAGuard g(device);
        auto dev = torch::Device(torch::kCUDA, device);
        auto img = torch::from_blob((void*)image_dev, {3, H, W}, torch::TensorOptions().dtype(torch::kFloat32).device(dev));
        auto mask = torch::from_blob((void*)mask_dev, {H, W}, torch::TensorOptions().dtype(mask_i64 ? torch::kInt64 : torch::kInt32).device(dev));
        SliceBox box;
        const torch::Tensor sl = slice_image(img, mask, label, box);
        if (!sl.defined()) return;
        found = 1;
        box_out[0] = box.x; box_out[1] = box.y; box_out[2] = box.width; box_out[3] = box.height;
        TORCH_CHECK(sl.numel() <= capacity, "slice buffer too small");
        torch::from_blob((void*)slice_out_dev, {sl.numel()}, torch::TensorOptions().dtype(torch::kUInt8).device(dev)).copy_(sl.reshape({-1}));
        torch::cuda::synchronize(device);
    });
    return rc ? rc : (found ? 0 : -1);
}

int icrh_reassemble(const char* image_path, const char* out_filename, int device, uint8_t* canvas_out_dev, int H, int W) {
    return guarded([&] {
        if (canvas_out_dev) {
            torch::Tensor c = reassemble_to_tensor(image_path, device);
            TORCH_CHECK(c.size(0) == H && c.size(1) == W, "canvas size differs from the caller's buffer");
            torch::from_blob(canvas_out_dev, {H, W, 4}, torch::TensorOptions().dtype(torch::kUInt8).device(torch::Device(torch::kCUDA, device))).copy_(c);
            torch::cuda::synchronize(device);
        }
        if (out_filename && !reassemble(image_path, out_filename)) throw std::runtime_error("reassemble failed");
    });
}

int icrh_write_metadata(const char* path, int n, const int32_t* label, const int32_t* x, const int32_t* y, const int32_t* w,
                        const int32_t* h, const char* names, uint32_t image_w, uint32_t image_h) {
    return guarded([&] {
        std::vector<SliceMetadata> m(n);
        const char* p = names;
        for (int i = 0; i < n; ++i) {
            m[i].label = label[i]; m[i].x = x[i]; m[i].y = y[i]; m[i].width = w[i]; m[i].height = h[i];
            m[i].filename = p;
            p += m[i].filename.size() + 1;
        }
        write_metadata_binary(m, path, image_w, image_h);
    });
}

int icrh_read_metadata_count(const char* path, uint32_t* image_w, uint32_t* image_h) {
    int n = -1;
    const int rc = guarded([&] { n = (int)read_metadata_binary(path, *image_w, *image_h).size(); });
    return rc ? -rc : n;
}

}
