// FeatureTrackerCuda.cpp -- see FeatureTrackerCuda.h.
#include "FeatureTrackerCuda.h"

#include <algorithm>
#include <random>
#include <stdexcept>

namespace vio_360 {

void FeatureTrackerCuda::check(int rc, const char* what) const {
    if (rc != VIO360_OK)
        throw std::runtime_error(std::string(what) + ": " + vio360_last_error(m_ctx));
}

FeatureTrackerCuda::FeatureTrackerCuda(int width, int height, int max_features, float min_distance,
                                       float quality_level, int boundary_margin, int device)
    : m_width(width), m_height(height), m_max_features(max_features), m_min_distance(min_distance),
      m_quality_level(quality_level), m_boundary_margin(boundary_margin) {
    if (vio360_create(device, &m_ctx) != VIO360_OK)
        throw std::runtime_error(std::string("FeatureTrackerCuda: ") + vio360_last_error(nullptr));
    ensure_capacity(max_features > 0 ? max_features : 1);
}

FeatureTrackerCuda::~FeatureTrackerCuda() {
    if (m_seq) vio360_seq_destroy(m_seq);
    delete m_seq_rng;
    vio360_destroy(m_ctx);
}

void FeatureTrackerCuda::RegisterFrameBuffer(const cv::Mat& image) {
    if (image.empty()) return;
    const void* p = image.data;
    const size_t bytes = (size_t)image.step * (size_t)image.rows;
    for (size_t i = 0; i < m_registered.size(); ++i)
        if (m_registered[i].first == p && m_registered[i].second >= bytes) {      // seen before: most recently used
            std::rotate(m_registered.begin() + i, m_registered.begin() + i + 1, m_registered.end());
            return;
        }
    check(vio360_host_register(m_ctx, p, bytes), "vio360_host_register");
    m_registered.emplace_back(p, bytes);
}

void FeatureTrackerCuda::UnregisterFrameBuffer(const cv::Mat& image) {
    for (size_t i = 0; i < m_registered.size(); ++i)
        if (m_registered[i].first == image.data) {
            check(vio360_host_unregister(m_ctx, image.data), "vio360_host_unregister");
            m_registered.erase(m_registered.begin() + i);
            return;
        }
}

void FeatureTrackerCuda::touch_frame(const cv::Mat& image) {
    if (!m_auto_register) return;
    RegisterFrameBuffer(image);
    while (m_registered.size() > m_max_registered) {
        check(vio360_host_unregister(m_ctx, m_registered.front().first), "vio360_host_unregister");
        m_registered.erase(m_registered.begin());
    }
}

void FeatureTrackerCuda::ResetSequence() {
    if (m_seq) { vio360_seq_destroy(m_seq); m_seq = nullptr; }
}

void FeatureTrackerCuda::TrackFeatures(const cv::Mat& image, FrameFeatures& out) {
    if (image.empty() || image.cols != m_width || image.rows != m_height)
        throw std::runtime_error("FeatureTrackerCuda::TrackFeatures: image must be 8-bit single channel of the configured size");
    touch_frame(image);
    ensure_capacity(2 * std::max(m_max_features, 1));      // may re-configure: only before the sequence exists
    if (!m_seq) {
        vio360_seq_config cfg;
        cfg.max_features = m_max_features; cfg.min_distance = m_min_distance; cfg.quality_level = m_quality_level;
        cfg.grid_cols = m_grid_cols; cfg.grid_rows = m_grid_rows; cfg.max_features_per_grid = m_grid_cap;
        cfg.remove_clustered = m_cluster ? 1 : 0; cfg.clustered_std_ratio = m_cluster_ratio;
        cfg.ransac_iterations = m_ransac_max_iterations; cfg.ransac_threshold_deg = m_ransac_threshold;
        check(vio360_seq_create(m_ctx, &cfg, &m_seq), "vio360_seq_create");
    }
    if (!m_seq_rng) { std::random_device rd; m_seq_rng = new std::mt19937(m_ransac_seeded ? m_ransac_seed : rd()); }   // FeatureTracker.cpp:273-274
    std::vector<uint32_t> words((size_t)std::max(m_ransac_max_iterations, 1) * 3);
    for (uint32_t& w : words) w = (uint32_t)(*m_seq_rng)();
    const int cap = vio360_seq_capacity(m_seq);
    out.ids.resize((size_t)cap); out.points.resize((size_t)cap); out.track_counts.resize((size_t)cap);
    int n = 0;
    vio360_frame_job job;
    job.seq = m_seq; job.image = image.data; job.pitch = (size_t)image.step; job.ransac_words = words.data();
    job.ids = out.ids.data(); job.xy = reinterpret_cast<float*>(out.points.data()); job.track_count = out.track_counts.data();
    job.n_features = &n; job.n_tracked = &out.num_tracked; job.n_detected = &out.num_detected;
    check(vio360_fe_track_frame_batch(m_ctx, &job, 1, VIO360_MEM_HOST), "vio360_fe_track_frame_batch");
    out.ids.resize((size_t)n); out.points.resize((size_t)n); out.track_counts.resize((size_t)n);
    m_num_tracked = out.num_tracked; m_num_detected = out.num_detected;
}

void FeatureTrackerCuda::ensure_capacity(int n_points) {
    if (n_points <= m_point_capacity) return;
    ResetSequence();                 // vio360_fe_configure invalidates the device-resident sequence
    m_point_capacity = n_points;
    // FeatureTracker.cpp:33-35: window 21x21, maxLevel 3, TermCriteria(EPS|COUNT, 30, 0.01f); the float
    // epsilon widens to double and doubles as minEigThreshold (:240).  Polar ratio 0.15 (Camera.h:76,84).
    const double eps = (double)0.01f;
    check(vio360_fe_configure(m_ctx, m_width, m_height, 3, 21, 30, eps, eps, 0.15f, m_boundary_margin,
                              m_point_capacity, 1),
          "vio360_fe_configure");
}

std::vector<cv::Point2f> FeatureTrackerCuda::DetectNewFeatures(const cv::Mat& image, const cv::Mat& mask) {
    if (image.empty() || image.cols != m_width || image.rows != m_height)
        throw std::runtime_error("FeatureTrackerCuda::DetectNewFeatures: image must be 8-bit single channel of the configured size");
    std::vector<cv::Point2f> corners((size_t)(m_max_features > 0 ? m_max_features : 0));
    int n = 0;
    if (m_max_features <= 0) return corners;
    touch_frame(image);
    static_assert(sizeof(cv::Point2f) == 2 * sizeof(float), "Point2f must be two packed floats");
    check(vio360_fe_detect(m_ctx, image.data, (size_t)image.step, m_curr_key, mask.empty() ? nullptr : mask.data,
                           mask.empty() ? 0 : (size_t)mask.step, m_max_features, (double)m_quality_level,
                           (double)m_min_distance, reinterpret_cast<float*>(corners.data()), &n),
          "vio360_fe_detect");
    corners.resize((size_t)n);
    m_num_detected = n;
    return corners;
}

std::vector<cv::Point2f> FeatureTrackerCuda::DetectNewFeatures(const cv::Mat& image, const std::vector<cv::Point2f>& occupied) {
    ensure_capacity(1);
    touch_frame(image);
    std::vector<cv::Point2f> corners((size_t)std::max(m_max_features, 1));
    int n = 0;
    vio360_detect_job job;
    job.img = image.data; job.pitch = (size_t)image.step; job.key = m_curr_key;
    job.mask = nullptr; job.mask_pitch = 0;
    job.out_xy = reinterpret_cast<float*>(corners.data()); job.out_n = &n;
    job.occupied_xy = occupied.empty() ? nullptr : reinterpret_cast<const float*>(occupied.data());
    job.n_occupied = (int)occupied.size();
    job.occupied_radius = static_cast<int>(m_min_distance);   // FeatureTracker.cpp:398
    check(vio360_fe_detect_batch(m_ctx, &job, 1, m_max_features, (double)m_quality_level, (double)m_min_distance, VIO360_MEM_HOST),
          "vio360_fe_detect_batch");
    corners.resize((size_t)n);
    m_num_detected = n;
    return corners;
}

int FeatureTrackerCuda::TrackOpticalFlow(const cv::Mat& prev_image, const cv::Mat& curr_image,
                                         const std::vector<cv::Point2f>& prev_points,
                                         std::vector<cv::Point2f>& curr_points, std::vector<uchar>& status) {
    if (prev_points.empty()) return 0;   // FeatureTracker.cpp:233-235: outputs untouched
    const int n = (int)prev_points.size();
    ensure_capacity(n);
    touch_frame(prev_image); touch_frame(curr_image);
    curr_points.resize((size_t)n);
    status.resize((size_t)n);
    m_keep.resize((size_t)n);
    vio360_track_job job;
    job.prev = prev_image.data; job.prev_pitch = (size_t)prev_image.step; job.prev_key = m_prev_key;
    job.cur = curr_image.data; job.cur_pitch = (size_t)curr_image.step; job.cur_key = m_curr_key;
    job.prev_xy = reinterpret_cast<const float*>(prev_points.data()); job.n = n;
    job.cur_xy = reinterpret_cast<float*>(curr_points.data()); job.status = status.data();
    job.err = nullptr;      // the reference requests err and discards it (:237): not computed (status is unaffected)
    job.keep = m_keep.data();
    check(vio360_fe_track_batch(m_ctx, &job, 1, VIO360_MEM_HOST), "vio360_fe_track_batch");
    int num_tracked = 0;
    for (size_t i = 0; i < status.size(); ++i)
        if (status[i]) num_tracked++;
    m_num_tracked = num_tracked;
    return num_tracked;
}

int FeatureTrackerCuda::RejectOutliersRotationRANSAC(const std::vector<cv::Point2f>& prev_points,
                                                     const std::vector<cv::Point2f>& curr_points,
                                                     std::vector<uchar>& inlier_mask) {
    const int n = (int)prev_points.size();
    if (n < 3) {   // FeatureTracker.cpp:258-261
        inlier_mask.resize(prev_points.size(), 1);
        return n;
    }
    ensure_capacity(1);
    // the draw of FeatureTracker.cpp:273-289
    std::random_device rd;
    std::mt19937 gen(m_ransac_seeded ? m_ransac_seed : rd());
    std::uniform_int_distribution<> dis(0, n - 1);
    std::vector<int32_t> samples((size_t)m_ransac_max_iterations * 3);
    for (int iter = 0; iter < m_ransac_max_iterations; ++iter) {
        int32_t* idx = samples.data() + 3 * (size_t)iter;
        int have = 0;
        while (have < 3) {
            const int k = dis(gen);
            if (std::find(idx, idx + have, k) == idx + have) idx[have++] = k;
        }
    }
    inlier_mask.assign((size_t)n, 0);
    int best_inliers = 0;
    check(vio360_fe_ransac_rotation(m_ctx, reinterpret_cast<const float*>(prev_points.data()),
                                    reinterpret_cast<const float*>(curr_points.data()), n, samples.data(),
                                    m_ransac_max_iterations, m_ransac_threshold, inlier_mask.data(), &best_inliers),
          "vio360_fe_ransac_rotation");
    return best_inliers;
}

std::vector<uchar> FeatureTrackerCuda::RemoveClusteredFeatures(const std::vector<cv::Point2f>& points, float clustered_std_ratio,
                                                               int grid_cols, int grid_rows) {
    std::vector<uchar> keep(points.size(), 1);
    if (points.empty()) return keep;
    const int n = (int)points.size();
    check(vio360_fe_cluster_filter(m_ctx, reinterpret_cast<const float*>(points.data()), &n, 1, n, m_width, m_height, grid_cols,
                                   grid_rows, clustered_std_ratio, keep.data(), VIO360_MEM_HOST), "vio360_fe_cluster_filter");
    return keep;
}

std::vector<uchar> FeatureTrackerCuda::LimitFeaturesPerGrid(const std::vector<cv::Point2f>& points, const std::vector<int>& track_counts,
                                                            int max_features_per_grid, int grid_cols, int grid_rows) {
    if (track_counts.size() != points.size()) throw std::runtime_error("LimitFeaturesPerGrid: one track count per point");
    std::vector<uchar> keep(points.size(), 0);
    if (points.empty()) return keep;
    const int n = (int)points.size();
    check(vio360_fe_grid_cap(m_ctx, reinterpret_cast<const float*>(points.data()), track_counts.data(), &n, 1, n, m_width, m_height,
                             grid_cols, grid_rows, max_features_per_grid, keep.data(), VIO360_MEM_HOST), "vio360_fe_grid_cap");
    return keep;
}

}  // namespace vio_360
