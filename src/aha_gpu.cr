# aha_gpu.cr -- Crystal binding of libaha_b200.so for chenkovsky/aha (drop-in for the Aha::AC scan path).
#
# NOT COMPILED IN THIS REPOSITORY'S BUILD IMAGE (no `crystal` there): written against include/aha_b200.h,
# which is exercised through ctypes by tests/.  Usage in an application that already uses aha:
#
#     require "aha"
#     require "aha_gpu"
#     ac  = Aha::AC.compile %w(我 我是 是中)      # unchanged reference builder (ac.cr:62-112)
#     gpu = ac.to_gpu                              # flatten + upload
#     gpu.match("我是中国人") { |hit| ... }         # same hits, same order as ac.match (ac.cr:280-295)
#     gpu.match_batch(docs) { |doc_idx, hit| ... } # throughput form
#
@[Link("aha_b200")]
lib LibAhaB200
  struct Hit
    start : Int32
    end_ : Int32
    value : Int32
  end

  struct Hits
    hits : Hit*
    n_hits : Int64
    doc_hit_off : Int64*
    n_docs : Int64
    on_device : Int32
    device : Int32
    scan_ms : Float64
    walk_ms : Float64
    expand_ms : Float64
    opaque : Void*
  end

  UNITS_CHAR    =  1_u32
  SEP           =  2_u32
  LONGEST       =  4_u32
  INTERSECTABLE =  8_u32
  TRIE_LONGEST  = 16_u32
  IGNORE_CASE   = 32_u32
  SEP_CODEPOINT = 128_u32 # Array(Char) / Slice(Char) forms of match(seq, sep): neighbour test on code points (ac.cr:342-364)

  fun init = aha_b200_init(devices : Int32*, n : Int32) : Int32
  fun last_error = aha_b200_last_error : UInt8*
  fun dfa_from_cedar = aha_b200_dfa_from_cedar(nodes : Void*, array_size : Int64, fails : Int32*, output : Int32*,
                                               key_lens : UInt32*, n_keys : Int64, out : Void**) : Int32
  fun dfa_free = aha_b200_dfa_free(dfa : Void*) : Void
  fun scan = aha_b200_scan(dfa : Void*, text : UInt8*, doc_off : Int64*, n_docs : Int64, mode : UInt32,
                           sep_bits : UInt8*, sep_size : Int32, out : Hits**) : Int32
  fun hits_free = aha_b200_hits_free(hits : Hits*) : Void
  # text and results already on the device, no allocation, no host wait (no counterpart in the reference)
  fun scan_device_noalloc = aha_b200_scan_device_noalloc(dfa : Void*, dev_text : Void*, dev_doc_off : Void*, n_docs : Int64,
                                                         total_bytes : Int64, dev_hits : Void*, hit_cap : Int64,
                                                         dev_doc_hit_off : Void*, dev_n_hits : Void*, stream : Void*) : Int32
end

module Aha
  # GPU twin of Aha::AC: same method names and block protocol (ac.cr:280-340, 297-319; matcher.cr:34-46).
  class GpuAC
    def initialize(@dfa : Void*)
    end

    def finalize
      LibAhaB200.dfa_free(@dfa)
    end

    private def check(rc : Int32)
      raise String.new(LibAhaB200.last_error) if rc != 0 # the reference raises Strings (ac.cr:66, 322)
    end

    private def sep_bytes(sep : BitArray) : {Bytes, Int32}
      raise "sep BitArray size > 256 is not supported" if sep.size > 256 # ac.cr:322
      bits = Bytes.new(32, 0_u8)
      sep.each_with_index { |b, i| bits[i >> 3] |= (1_u8 << (i & 7)) if b }
      {bits, sep.size}
    end

    # documents back to back + offsets -> yields (doc_idx, Hit) in document order, reference emission order
    def scan_batch(docs : Array(Bytes), mode : UInt32, sep : BitArray? = nil, &block)
      offs = Array(Int64).new(docs.size + 1, 0_i64)
      docs.each_with_index { |d, i| offs[i + 1] = offs[i] + d.size }
      text = Bytes.new(offs[-1])
      docs.each_with_index { |d, i| d.copy_to(text + offs[i]) }
      bits, size = sep ? sep_bytes(sep) : {Bytes.new(32, 0_u8), 0}
      mode |= LibAhaB200::SEP if sep
      res = Pointer(LibAhaB200::Hits).null
      check LibAhaB200.scan(@dfa, text.to_unsafe, offs.to_unsafe, docs.size.to_i64, mode, bits.to_unsafe, size, pointerof(res))
      begin
        r = res.value
        docs.size.times do |d|
          (r.doc_hit_off[d]...r.doc_hit_off[d + 1]).each do |i|
            h = r.hits[i]
            yield d, Aha::Hit.new(h.start, h.end_, h.value)
          end
        end
      ensure
        LibAhaB200.hits_free(res)
      end
    end

    def match_batch(docs : Array(Bytes), &block)
      scan_batch(docs, 0_u32) { |d, h| yield d, h }
    end

    # Bytes overloads: byte offsets (ac.cr:280-286, 321-340)
    def match(seq : Bytes, &block)
      scan_batch([seq], 0_u32) { |_, h| yield h }
    end

    def match(seq : Bytes, sep : BitArray, &block)
      scan_batch([seq], 0_u32, sep) { |_, h| yield h }
    end

    # String overloads: character offsets (matcher.cr:34-46)
    def match(seq : String, &block)
      scan_batch([seq.to_slice], LibAhaB200::UNITS_CHAR) { |_, h| yield h }
    end

    def match(seq : String, sep : BitArray, &block)
      scan_batch([seq.to_slice], LibAhaB200::UNITS_CHAR, sep) { |_, h| yield h }
    end

    # Array(Char) / Slice(Char) overloads: character offsets (ac.cr:288-295); with `sep` the neighbour test is made on
    # the neighbouring character's code point (ac.cr:342-364), which the library does under SEP_CODEPOINT
    def match(seq : Array(Char) | Slice(Char), &block)
      scan_batch([String.build { |sb| seq.each { |c| sb << c } }.to_slice], LibAhaB200::UNITS_CHAR) { |_, h| yield h }
    end

    def match(seq : Array(Char) | Slice(Char), sep : BitArray, &block)
      m = LibAhaB200::UNITS_CHAR | LibAhaB200::SEP_CODEPOINT
      scan_batch([String.build { |sb| seq.each { |c| sb << c } }.to_slice], m, sep) { |_, h| yield h }
    end

    # ac.cr:312-319
    def match_longest(seq : Array(Char) | Slice(Char), intersectable = false, &block)
      m = LibAhaB200::LONGEST | LibAhaB200::UNITS_CHAR | (intersectable ? LibAhaB200::INTERSECTABLE : 0_u32)
      scan_batch([String.build { |sb| seq.each { |c| sb << c } }.to_slice], m) { |_, h| yield h }
    end

    # ac.cr:297-310
    def match_longest(seq : Bytes, intersectable = false, &block)
      m = LibAhaB200::LONGEST | (intersectable ? LibAhaB200::INTERSECTABLE : 0_u32)
      scan_batch([seq], m) { |_, h| yield h }
    end

    def match_longest(seq : String, intersectable = false, &block)
      m = LibAhaB200::LONGEST | LibAhaB200::UNITS_CHAR | (intersectable ? LibAhaB200::INTERSECTABLE : 0_u32)
      scan_batch([seq.to_slice], m) { |_, h| yield h }
    end

    # CedarX#match_longest(String, ignore_case, sep) (cedar.cr:1075-1121); sep restricted to code points < 256
    def trie_match_longest(seq : String, ignore_case = false, sep : BitArray? = nil, &block)
      m = LibAhaB200::TRIE_LONGEST | (ignore_case ? LibAhaB200::IGNORE_CASE : 0_u32)
      scan_batch([seq.to_slice], m, sep) { |_, h| yield h }
    end

    # sep : Array(Char) (cedar.cr:1071-1073): the kernel's separator set has 256 bits, wider characters are refused
    def trie_match_longest(seq : String, ignore_case : Bool, sep : Array(Char) | Slice(Char), &block)
      raise "aha_gpu: separators must be code points below U+0100" if sep.any? { |c| c.ord >= 256 }
      bits = BitArray.new(256)
      sep.each { |c| bits[c.ord] = true }
      trie_match_longest(seq, ignore_case, bits) { |h| yield h }
    end

    # CedarX#gsub(String, ignore_case, sep) { |hit| replacement } (cedar.cr:1123-1177): the reference loop is the
    # match_longest loop with `sb <<` on every branch, so the string is rebuilt from the GPU's hit stream.
    def trie_gsub(seq : String, ignore_case = false, sep : BitArray? = nil, &block)
      chars = seq.chars
      String.build do |sb|
        pos = 0
        trie_match_longest(seq, ignore_case, sep) do |h|
          (pos...h.start).each { |i| sb << chars[i] }
          sb << (yield h)
          pos = h.end
        end
        (pos...chars.size).each { |i| sb << chars[i] }
      end
    end
  end

  # Reopened: the ivars of ac.cr:35-39 and the protected getters of cedar.cr:147-148 are reachable from inside.
  class ACX(T)
    def to_gpu : GpuAC
      {% raise "aha_gpu supports Aha::AC (Int32) only" unless T == Int32 %}
      dfa = Pointer(Void).null
      rc = LibAhaB200.dfa_from_cedar(@da.gpu_nodes, @da.gpu_array_size.to_i64, @fails.ptr, @output.ptr.as(Int32*),
        @key_lens.ptr, @key_lens.size, pointerof(dfa))
      raise String.new(LibAhaB200.last_error) if rc != 0
      GpuAC.new(dfa)
    end
  end

  class CedarX(T)
    # A trie alone on the GPU (for trie_match_longest / trie_gsub, cedar.cr:1075-1177): the flattener wants the arrays of
    # a compiled automaton, so the trie goes through ACX.compile(da) (ac.cr:71-112), which leaves the trie itself untouched.
    def to_gpu : GpuAC
      ACX(T).compile(self).to_gpu
    end

    # raw views for the flattener: Node(Int32)[array_size], 12 bytes each (cedar.cr:29-36)
    def gpu_nodes : Void*
      @array.as(Void*)
    end

    def gpu_array_size
      @array_size
    end
  end
end
