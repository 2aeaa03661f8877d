struct RandomTEA {
private:
  // Tiny Encryption Algorithm (TEA) to calculate a the seed per launch index and iteration.
  // This results in a ton of integer instructions! Use the smallest N necessary.
  template<unsigned int N>
  __forceinline__ __device__ void tea(unsigned int& _v0, unsigned int& _v1)
  {
    unsigned int v0 = _v0; // Operate on registers to avoid slow down!
    unsigned int v1 = _v1;
    unsigned int sum = 0;

    for (int i = 0; i < N; i++) {
      sum += 0x9e3779b9;
      v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
      v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }

    _v0 = v0;
    _v1 = v1;
  }

private:
  unsigned int v0, v1;

public:
  __forceinline__ __device__ RandomTEA(const unsigned int idx, const unsigned int seed)
  {
    this->v0 = idx;
    this->v1 = seed;
  }
    
  __forceinline__ __device__ float get_float()
  {
    return get_floats().x;
  }

  __forceinline__ __device__ float2 get_floats()
  {
    tea<16>(this->v0, this->v1);
    const float tofloat = 2.3283064365386962890625e-10f; // 1/2^32
    return make_float2(this->v0 * tofloat, this->v1 * tofloat);
  }
};
