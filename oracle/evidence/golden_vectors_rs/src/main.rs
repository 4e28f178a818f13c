//! Golden-vector check against the reference's own crates: recomputes every hash of
//! tests/golden/ahash_vectors.json with `datafusion::common::hash_utils::create_hashes` and
//! `ahash::RandomState::with_seeds(0, 0, 0, 0)` - exactly what `BatchPartitioner` does for `Partitioning::Hash`
//! (call site: ballista/core/src/execution_plans/shuffle_writer.rs:336-346) - and reports every value that differs from
//! the one the oracle / the CUDA kernels are held to. Exit code 0 = parity of the hasher AND of DataFusion's glue around
//! it (seed derivation, which Rust type each Arrow type is hashed as, combine_hashes, dictionary handling) is pinned.
use std::sync::Arc;

use ahash::RandomState;
use datafusion::arrow::array::*;
use datafusion::arrow::datatypes::{DataType, Int32Type};
use datafusion::common::hash_utils::create_hashes;
use serde_json::Value;

fn hashes(cols: &[ArrayRef]) -> Vec<u64> {
    let random_state = RandomState::with_seeds(0, 0, 0, 0);
    let mut buf = vec![0u64; cols[0].len()];
    create_hashes(cols, &random_state, &mut buf).unwrap();
    buf
}

fn unhex(s: &str) -> Vec<u8> {
    (0..s.len()).step_by(2).map(|i| u8::from_str_radix(&s[i..i + 2], 16).unwrap()).collect()
}

fn main() {
    let path = std::env::args().nth(1).expect("path to tests/golden/ahash_vectors.json");
    let v: Value = serde_json::from_str(&std::fs::read_to_string(path).unwrap()).unwrap();
    let mut bad = 0usize;
    let mut check = |name: &str, cols: Vec<ArrayRef>, want: Vec<u64>| {
        let got = hashes(&cols);
        for (i, (g, w)) in got.iter().zip(want.iter()).enumerate() {
            if g != w {
                bad += 1;
                println!("MISMATCH {name}[{i}]: reference {g} != golden {w}");
            }
        }
        println!("{name}: {} values checked", got.len());
    };
    let rows = |k: &str| v[k].as_array().unwrap().clone();
    let last = |r: &Value| r.as_array().unwrap().last().unwrap().as_u64().unwrap();

    let r = rows("i64");
    check("i64", vec![Arc::new(Int64Array::from_iter_values(r.iter().map(|x| x[0].as_i64().unwrap())))], r.iter().map(last).collect());
    let r = rows("i32");
    check("i32", vec![Arc::new(Int32Array::from_iter_values(r.iter().map(|x| x[0].as_i64().unwrap() as i32)))], r.iter().map(last).collect());
    let r = rows("u32");
    check("u32", vec![Arc::new(UInt32Array::from_iter_values(r.iter().map(|x| x[0].as_u64().unwrap() as u32)))], r.iter().map(last).collect());
    let r = rows("i16");
    check("i16", vec![Arc::new(Int16Array::from_iter_values(r.iter().map(|x| x[0].as_i64().unwrap() as i16)))], r.iter().map(last).collect());
    let r = rows("i8");
    check("i8", vec![Arc::new(Int8Array::from_iter_values(r.iter().map(|x| x[0].as_i64().unwrap() as i8)))], r.iter().map(last).collect());
    let r = rows("f64");
    check("f64", vec![Arc::new(Float64Array::from_iter_values(r.iter().map(|x| x[0].as_f64().unwrap())))], r.iter().map(last).collect());
    let r = rows("bool");
    check("bool", vec![Arc::new(BooleanArray::from(r.iter().map(|x| x[0].as_i64().unwrap() != 0).collect::<Vec<bool>>()))], r.iter().map(last).collect());
    let r = rows("utf8");
    let strings: Vec<String> = r.iter().map(|x| x[0].as_str().unwrap().to_string()).collect();
    check("utf8", vec![Arc::new(StringArray::from(strings.clone()))], r.iter().map(last).collect());
    // dictionary-encoded strings hash like plain strings
    let dict: DictionaryArray<Int32Type> = strings.iter().map(|s| s.as_str()).collect();
    check("dict_utf8", vec![Arc::new(dict)], r.iter().map(last).collect());
    let r = rows("binary");
    let bytes: Vec<Vec<u8>> = r.iter().map(|x| unhex(x[0].as_str().unwrap())).collect();
    check("binary", vec![Arc::new(BinaryArray::from_iter_values(bytes.iter().map(|b| b.as_slice())))], r.iter().map(last).collect());
    let r = rows("decimal128");
    let dec = Decimal128Array::from_iter_values(r.iter().map(|x| x[0].as_str().unwrap().parse::<i128>().unwrap()))
        .with_data_type(DataType::Decimal128(38, 0));
    check("decimal128", vec![Arc::new(dec)], r.iter().map(last).collect());
    let r = rows("i64_utf8");
    check(
        "i64_utf8",
        vec![
            Arc::new(Int64Array::from_iter_values(r.iter().map(|x| x[0].as_i64().unwrap()))),
            Arc::new(StringArray::from(r.iter().map(|x| x[1].as_str().unwrap().to_string()).collect::<Vec<String>>())),
        ],
        r.iter().map(last).collect(),
    );
    if bad > 0 {
        println!("{bad} mismatches: the restatement differs from the reference build");
        std::process::exit(1);
    }
    println!("all golden vectors reproduced by the reference's crates");
}
