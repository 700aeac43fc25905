// TEST INFRASTRUCTURE ONLY (oracle build). Brute-force stand-in for the subset of flann 1.9.x
// that /root/reference/planner/src/utils/tree.h:44,63,79-84,104-109,142,166,174 uses (flann is
// an un-vendored system dependency of the reference and is absent from this image).
// Semantics restated from flann's published behaviour: Index<L2_Simple<T>> with
// KDTreeSingleIndexParams and default SearchParams performs an EXACT k-NN search under the
// squared Euclidean distance; results are ordered by increasing distance.  Tie order between
// equidistant points is not pinned by any reference test; this stand-in breaks ties by lowest
// insertion index (documented choice, mirrored by the CUDA kNN kernel).
// flann keeps a pointer to the caller's point storage (it does not copy unless asked), which
// the reference relies on (tree.h:63,142); the stand-in does the same.
#pragma once
#include <vector>
#include <cstddef>
#include <algorithm>

namespace flann {

template <typename T>
class Matrix {
public:
	Matrix() : rows(0), cols(0), m_data(nullptr) { }
	Matrix(T* data, size_t rows_, size_t cols_) : rows(rows_), cols(cols_), m_data(data) { }
	T* operator[](size_t i) const { return m_data + i * cols; }
	T* ptr() const { return m_data; }

public:
	size_t rows, cols;

private:
	T* m_data;
};

template <typename T>
struct L2_Simple {
	typedef T ElementType;
	typedef T ResultType;
};

struct KDTreeSingleIndexParams {
	KDTreeSingleIndexParams(int /*leaf_max_size*/ = 10, bool /*reorder*/ = true) { }
};

struct SearchParams {
	SearchParams(int /*checks*/ = 32, float /*eps*/ = 0.0f, bool /*sorted*/ = true) { }
};

template <typename Distance>
class Index {
public:
	typedef typename Distance::ElementType ElementType;
	typedef typename Distance::ResultType DistanceType;

	Index(const KDTreeSingleIndexParams& = KDTreeSingleIndexParams()) { }

	void buildIndex(const Matrix<ElementType>& points)
	{
		m_points.clear();
		m_dim = points.cols;
		addPoints(points);
	}

	void addPoints(const Matrix<ElementType>& points, float /*rebuild_threshold*/ = 2)
	{
		if (m_dim == 0)
			m_dim = points.cols;
		for (size_t i = 0; i < points.rows; i++)
			m_points.push_back(points[i]);
	}

	ElementType* getPoint(size_t i) { return m_points[i]; }
	size_t size() const { return m_points.size(); }

	int knnSearch(const Matrix<ElementType>& queries, Matrix<int>& indices, Matrix<DistanceType>& dists, size_t knn, const SearchParams&) const
	{
		int count = 0;
		for (size_t q = 0; q < queries.rows; q++) {
			std::vector<std::pair<DistanceType, int>> all;
			all.reserve(m_points.size());
			for (size_t p = 0; p < m_points.size(); p++) {
				DistanceType d = 0;
				for (size_t k = 0; k < m_dim; k++) {
					DistanceType diff = queries[q][k] - m_points[p][k];
					d += diff * diff;
				}
				all.push_back({ d, (int)p });
			}
			size_t n = std::min(knn, all.size());
			std::partial_sort(all.begin(), all.begin() + n, all.end());
			for (size_t k = 0; k < n; k++) {
				indices[q][k] = all[k].second;
				dists[q][k] = all[k].first;
			}
			count += (int)n;
		}
		return count;
	}

private:
	std::vector<ElementType*> m_points;
	size_t m_dim = 0;
};

}
