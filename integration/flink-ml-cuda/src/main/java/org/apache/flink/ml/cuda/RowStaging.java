/*
 * Host staging for micro-batches on their way to the GPU: a direct, little-endian ByteBuffer of rows
 * (row-major doubles), the layout fmlk_append_points / fmlk_predict / fmlk_online_round read in place.
 * NOT compiled in this repository (no JDK in the build environment); see INTEGRATION.md.
 */
package org.apache.flink.ml.cuda;

import org.apache.flink.ml.linalg.DenseVector;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;
import java.nio.DoubleBuffer;

final class RowStaging {
    private final int capacityRows;
    private ByteBuffer bytes;
    private DoubleBuffer doubles;
    private int dim = -1;
    private int rows = 0;

    RowStaging(int capacityRows) {
        this.capacityRows = capacityRows;
    }

    /** Copies the vector (object reuse is on in Flink ML jobs: the caller's array may change after we return). */
    void put(DenseVector v) {
        if (dim < 0) {
            dim = v.size();
            bytes = ByteBuffer.allocateDirect(capacityRows * dim * Double.BYTES).order(ByteOrder.LITTLE_ENDIAN);
            doubles = bytes.asDoubleBuffer();
        }
        if (v.size() != dim) {
            throw new IllegalArgumentException("Vector size mismatched."); // BLAS.java:36,70
        }
        doubles.put(v.values);
        rows++;
    }

    boolean full() { return rows == capacityRows; }
    boolean isEmpty() { return rows == 0; }
    int rows() { return rows; }
    int dim() { return dim; }
    ByteBuffer buffer() { return bytes; }

    void clear() {
        rows = 0;
        if (doubles != null) {
            doubles.clear();
        }
    }
}
